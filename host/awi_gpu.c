/*
 * awi_gpu.c — drop-in for the reference's awi.c codec drivers on top of gpu_codec.h.
 *
 * compress_to_awi / decompress_awi keep the reference's signatures, call order and
 * messages (awi.c:65-141, 153-230).  What changes is the loop body: instead of
 * transforming and bit-writing one macroblock at a time (awi.c:94-128, 184-217) the frame
 * is materialised once, handed to the GPU library in one call, and the payload goes to
 * bo->write in one call (the vtable accepts any size: bitrw.h:19, awi.c:25-33).
 * read_block / write_block are still invoked in strict raster order, which the
 * reference's BMP contexts rely on (bmp.c:61-63, 233-240).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "awi_host.h"
#include "../include/gpu_codec.h"

static int bmp_stride(int w) { return ((24 * w + 31) / 32) * 4; }      /* bmp.c:36 */

static bool ensure_gpu(void) {
    static int state = 0;                       /* 0 untried, 1 ok, -1 failed */
    if (state == 0) state = gpu_codec_init(1) ? 1 : -1;
    return state == 1;
}

static bool write_header(BitsOutput* bo, int width, int height) {      /* awi.c:75-83 */
    AWIHeader h = { { 'A', 'W', 'E', 'S', 'O', 'M', 'E', 'I' }, (uint32_t) width, (uint32_t) height };
    if (!bo->write(bo, &h, sizeof(h))) {
        fprintf(stderr, "Writing header error!\n");
        return false;
    }
    return true;
}

bool compress_rows_to_awi(const uint8_t* rows, int width, int height, int stride, BitsOutput* bo) {
    if (!ensure_gpu()) return false;
    if (!write_header(bo, width, height)) return false;
    uint8_t* payload = NULL;
    size_t n = 0;
    if (!gpu_codec_encode_bgr(rows, width, height, stride, &payload, &n)) {
        fprintf(stderr, "Encoding and writing error!\n");
        return false;
    }
    bool ok = n == 0 || bo->write(bo, payload, n);
    gpu_codec_free(payload);
    if (!ok) fprintf(stderr, "Encoding and writing error!\n");
    return ok;
}

bool compress_to_awi(PixelInput* pi, BitsOutput* bo) {
    if (pi->begin != NULL && !pi->begin(pi)) {
        fprintf(stderr, "Compressing: Preprocessing error!\n");
        return false;
    }
    const int width = pi->width, height = pi->height;
    const int stride = bmp_stride(width);
    const int bw = (width + 15) / 16, bh = (height + 15) / 16;
    uint8_t* rows = (uint8_t*) calloc((size_t) stride * height, 1);
    if (!rows) return false;
    RGBBlock16 blk;
    for (int by = 0; by < bh; by++) {
        for (int bx = 0; bx < bw; bx++) {
            if (!pi->read_block(pi, bx, by, &blk)) {
                fprintf(stderr, "Reading RGB blocks error!\n");
                free(rows);
                return false;
            }
            for (int dy = 0; dy < 16 && by * 16 + dy < height; dy++) {
                uint8_t* row = rows + (size_t) (height - 1 - (by * 16 + dy)) * stride;
                for (int dx = 0; dx < 16 && bx * 16 + dx < width; dx++) {
                    uint8_t* p = row + 3 * (bx * 16 + dx);
                    p[0] = blk.b.b[dy][dx]; p[1] = blk.g.b[dy][dx]; p[2] = blk.r.b[dy][dx];
                }
            }
        }
    }
    bool ok = compress_rows_to_awi(rows, width, height, stride, bo);
    free(rows);
    if (!ok) return false;
    if (pi->end != NULL && !pi->end(pi)) {
        fprintf(stderr, "Compressing: Postprocessing error!\n");
        return false;
    }
    return true;
}

bool decompress_awi_to_rows(const uint8_t* awi, size_t awi_len, int* width, int* height,
                            uint8_t** pixels, int* row_stride, int rgb24) {
    static const char magic[8] = { 'A', 'W', 'E', 'S', 'O', 'M', 'E', 'I' };
    if (awi_len < sizeof(AWIHeader)) { fprintf(stderr, "Reading header error!\n"); return false; }
    if (memcmp(awi, magic, 8) != 0) { fprintf(stderr, "Invalid file format!\n"); return false; }
    AWIHeader h;
    memcpy(&h, awi, sizeof(h));
    if (!ensure_gpu()) return false;
    const int w = (int) h.width, hgt = (int) h.height;
    const int stride = rgb24 ? 3 * w : bmp_stride(w);
    uint8_t* px = (uint8_t*) calloc((size_t) stride * hgt, 1);
    if (!px) return false;
    bool ok = rgb24 ? gpu_codec_decode_to_rgb24(awi + 16, awi_len - 16, w, hgt, px)
                    : gpu_codec_decode_to_bgr(awi + 16, awi_len - 16, w, hgt, px, stride);
    if (!ok) {
        fprintf(stderr, "Reading and decoding error!\n");
        free(px);
        return false;
    }
    *width = w; *height = hgt; *pixels = px; *row_stride = stride;
    return true;
}

bool decompress_awi(BitsInput* bi, PixelOutput* po) {
    AWIHeader header;
    if (!bi->read(bi, &header, sizeof(AWIHeader))) {
        fprintf(stderr, "Reading header error!\n");
        return false;
    }
    /* the payload length is not in the container: read until the source runs dry */
    size_t cap = 1 << 20, n = sizeof(header);
    uint8_t* buf = (uint8_t*) malloc(cap);
    if (!buf) return false;
    memcpy(buf, &header, sizeof(header));
    for (;;) {
        if (n == cap) {
            uint8_t* nb = (uint8_t*) realloc(buf, cap *= 2);
            if (!nb) { free(buf); return false; }
            buf = nb;
        }
        if (!bi->read(bi, buf + n, 1)) break;
        n++;
    }
    int w, h, stride;
    uint8_t* rgb = NULL;
    if (!decompress_awi_to_rows(buf, n, &w, &h, &rgb, &stride, 1)) { free(buf); return false; }
    free(buf);
    if (po->begin != NULL && !po->begin(po, w, h)) {
        fprintf(stderr, "Decompressing: Preprocessing error!\n");
        free(rgb);
        return false;
    }
    RGBBlock16 blk;
    for (int by = 0; by < (h + 15) / 16; by++) {
        for (int bx = 0; bx < (w + 15) / 16; bx++) {
            memset(&blk, 0, sizeof(blk));
            for (int dy = 0; dy < 16 && by * 16 + dy < h; dy++)
                for (int dx = 0; dx < 16 && bx * 16 + dx < w; dx++) {
                    const uint8_t* p = rgb + ((size_t) (by * 16 + dy) * w + bx * 16 + dx) * 3;
                    blk.r.b[dy][dx] = p[0]; blk.g.b[dy][dx] = p[1]; blk.b.b[dy][dx] = p[2];
                }
            if (!po->write_block(po, bx, by, &blk)) {
                fprintf(stderr, "Writing rgb blocks error!\n");
                free(rgb);
                return false;
            }
        }
    }
    free(rgb);
    if (po->end != NULL && !po->end(po)) {
        fprintf(stderr, "Decompressing: Postprocessing error!\n");
        return false;
    }
    return true;
}
