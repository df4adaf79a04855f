/*
 * oracle/ref_shim.c — TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Thin memory-backed harness around the UNMODIFIED reference sources, which are
 * compiled where they lie under /root/reference/src by oracle/Makefile into
 * oracle/_ref/libawi_ref.so.  Nothing of the reference is copied here: this file
 * only *calls* the reference's public functions
 *   compress_to_awi / decompress_awi           (awi.h:45-46)
 *   init_bmp_pixel_input / init_bmp_pixel_output / init_rgb_pixel_output (bmp.h:57-66)
 *   dct_block, quantize_block, zigzag_block, in_* (dct.h:12-22)
 *   rle_encode_chunk_and_write, read_and_rle_decode_chunk (rle.h:29-30)
 * through the reference's own plug-in vtables (BitsOutput.write / BitsInput.read,
 * bitrw.h:19,41) with a memory sink/source instead of fwrite/fread, and feeds
 * the reference's own BMP reader/writer from fmemopen()ed FILE*s so no disk I/O
 * is timed.
 */
#define _GNU_SOURCE
#include <stdio.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stdbool.h>

#include "awi.h"
#include "bmp.h"
#include "dct.h"
#include "rle.h"
#include "color.h"

/* ---------------------------------------------------------------- sinks */
typedef struct { uint8_t* p; size_t len, cap; bool overflow; } MemSink;
typedef struct { const uint8_t* p; size_t pos, len; } MemSource;

static bool sink_write(BitsOutput* self, const void* data, size_t size) {
    MemSink* s = (MemSink*) self->context;
    if (s->len + size > s->cap) { s->overflow = true; return false; }
    memcpy(s->p + s->len, data, size);
    s->len += size;
    return true;
}

static bool source_read(BitsInput* self, void* buf, size_t size) {
    MemSource* s = (MemSource*) self->context;
    if (s->pos + size > s->len) return false;
    memcpy(buf, s->p + s->pos, size);
    s->pos += size;
    return true;
}

static void sink_init(BitsOutput* bo, MemSink* s, uint8_t* out, size_t cap) {
    s->p = out; s->len = 0; s->cap = cap; s->overflow = false;
    bo->context = s; bo->write = &sink_write; bo->buffer = 0; bo->bit_count = 0;
}

static void source_init(BitsInput* bi, MemSource* s, const uint8_t* in, size_t n) {
    s->p = in; s->pos = 0; s->len = n;
    bi->context = s; bi->read = &source_read; bi->buffer = 0; bi->bit_count = 0;
}

/* ------------------------------------------------------ whole-file paths */

/* BMP file image in memory -> .awi file image (16-byte header + payload).
 * Returns the .awi length, 0 on failure.  Same calls as main.c:104-116. */
size_t ref_encode_bmp_mem(const uint8_t* bmp, size_t bmp_len, uint8_t* out, size_t cap) {
    FILE* fp = fmemopen((void*) bmp, bmp_len, "rb");
    if (!fp) return 0;
    PixelInput pi;
    BMPInputContext ctx = { .fp = fp };
    init_bmp_pixel_input(&pi, &ctx);
    BitsOutput bo; MemSink sink;
    sink_init(&bo, &sink, out, cap);
    bool ok = compress_to_awi(&pi, &bo);
    fclose(fp);
    return ok ? sink.len : 0;
}

/* .awi file image -> BMP file image (54-byte header, bottom-up BGR).
 * Same calls as main.c:137-149.  Returns BMP length, 0 on failure. */
size_t ref_decode_awi_to_bmp_mem(const uint8_t* awi, size_t n, uint8_t* out, size_t cap) {
    /* fmemopen() appends a NUL at the stream position on close, so the stream gets a
     * private buffer one byte larger than the file; zeroed because the reference's
     * row padding is otherwise indeterminate (bmp.c:154). */
    uint8_t* tmp = (uint8_t*) calloc(cap + 1, 1);
    if (!tmp) return 0;
    FILE* fp = fmemopen(tmp, cap + 1, "w+b");
    if (!fp) { free(tmp); return 0; }
    BitsInput bi; MemSource src;
    source_init(&bi, &src, awi, n);
    PixelOutput po;
    BMPOutputContext ctx = { .fp = fp };
    init_bmp_pixel_output(&po, &ctx);
    bool ok = decompress_awi(&bi, &po);
    size_t len = 0;
    if (ok) len = 54 + (size_t) ctx.row_size * (size_t) po.height;
    fclose(fp);
    if (ok && len <= cap) memcpy(out, tmp, len); else ok = false;
    free(tmp);
    return ok ? len : 0;
}

/* .awi file image -> packed top-down RGB24 (the `show` sink, main.c:162-181).
 * `rgb` must hold w*h*3 bytes; returns 1 on success and fills w,h. */
int ref_decode_awi_to_rgb_mem(const uint8_t* awi, size_t n, uint8_t* rgb, size_t cap, int* w, int* h) {
    BitsInput bi; MemSource src;
    source_init(&bi, &src, awi, n);
    PixelOutput po; TextureOutputContext tctx; uint8_t* pixels = NULL;
    init_rgb_pixel_output(&po, &tctx, &pixels);
    bool ok = decompress_awi(&bi, &po);
    if (ok) {
        size_t need = (size_t) po.width * po.height * 3;
        if (need <= cap) memcpy(rgb, pixels, need); else ok = false;
        *w = po.width; *h = po.height;
    }
    free(pixels);
    return ok ? 1 : 0;
}

/* ------------------------------------------------------- per-stage hooks */

/* awi.c:57-62 transform_block is static there; this makes the same three calls. */
void ref_transform_block(const uint8_t in[64], int qm, int chunk_out[64]) {
    Block_u8 b; Block_int d, q; Chunk c;
    memcpy(b.b, in, 64);
    dct_block(&b, &d);
    quantize_block(&d, &q, qm ? QM_CHROM : QM_LUMA);
    zigzag_block(&q, &c);
    memcpy(chunk_out, c.c, sizeof(c.c));
}

void ref_dct_block(const uint8_t in[64], int out[64]) {
    Block_u8 b; Block_int d;
    memcpy(b.b, in, 64);
    dct_block(&b, &d);
    memcpy(out, d.b, sizeof(d.b));
}

void ref_quantize_block(const int in[64], int qm, int out[64]) {
    Block_int a, q;
    memcpy(a.b, in, sizeof(a.b));
    quantize_block(&a, &q, qm ? QM_CHROM : QM_LUMA);
    memcpy(out, q.b, sizeof(q.b));
}

/* awi.c:146-151 inverse_transform_block */
void ref_inverse_transform_block(const int chunk[64], int qm, uint8_t out[64]) {
    Chunk c; Block_int z, q; Block_u8 o;
    memcpy(c.c, chunk, sizeof(c.c));
    in_zigzag_block(&c, &z);
    in_quantize_block(&z, &q, qm ? QM_CHROM : QM_LUMA);
    in_dct_block(&q, &o);
    memcpy(out, o.b, 64);
}

int ref_rle_encode_chunk(const int chunk[64], int prev_dc, uint8_t* out, size_t cap) {
    BitsOutput bo; MemSink sink; Chunk c;
    memcpy(c.c, chunk, sizeof(c.c));
    sink_init(&bo, &sink, out, cap);
    if (!rle_encode_chunk_and_write(&bo, &c, prev_dc)) return -1;
    return (int) sink.len;
}

/* returns bytes consumed, -1 on failure */
int ref_rle_decode_chunk(const uint8_t* in, size_t n, int prev_dc, int chunk_out[64]) {
    BitsInput bi; MemSource src; Chunk c;
    source_init(&bi, &src, in, n);
    if (!read_and_rle_decode_chunk(&bi, &c, prev_dc)) return -1;
    memcpy(chunk_out, c.c, sizeof(c.c));
    return (int) src.pos;
}

/* colour, one pixel at a time through the 16x16 block functions (color.c:14-67) */
void ref_rgb_to_ycbcr_px(const uint8_t* rgb, size_t npx, uint8_t* ycc) {
    RGBBlock16 in; YCbCrBlock_full out;
    for (size_t base = 0; base < npx; base += 256) {
        size_t m = npx - base < 256 ? npx - base : 256;
        memset(&in, 0, sizeof(in));
        for (size_t i = 0; i < m; i++) {
            in.r.b[i / 16][i % 16] = rgb[3 * (base + i) + 0];
            in.g.b[i / 16][i % 16] = rgb[3 * (base + i) + 1];
            in.b.b[i / 16][i % 16] = rgb[3 * (base + i) + 2];
        }
        rgb_block16_to_ycbcr(&in, &out);
        for (size_t i = 0; i < m; i++) {
            int py = i / 16, px = i % 16;
            ycc[3 * (base + i) + 0] = out.y[py / 8][px / 8].b[py % 8][px % 8];
            ycc[3 * (base + i) + 1] = out.cb.b[py][px];
            ycc[3 * (base + i) + 2] = out.cr.b[py][px];
        }
    }
}

void ref_ycbcr_to_rgb_px(const uint8_t* ycc, size_t npx, uint8_t* rgb) {
    YCbCrBlock_full in; RGBBlock16 out;
    for (size_t base = 0; base < npx; base += 256) {
        size_t m = npx - base < 256 ? npx - base : 256;
        memset(&in, 0, sizeof(in));
        for (size_t i = 0; i < m; i++) {
            int py = i / 16, px = i % 16;
            in.y[py / 8][px / 8].b[py % 8][px % 8] = ycc[3 * (base + i) + 0];
            in.cb.b[py][px] = ycc[3 * (base + i) + 1];
            in.cr.b[py][px] = ycc[3 * (base + i) + 2];
        }
        ycbcr_block_to_rgb(&in, &out);
        for (size_t i = 0; i < m; i++) {
            rgb[3 * (base + i) + 0] = out.r.b[i / 16][i % 16];
            rgb[3 * (base + i) + 1] = out.g.b[i / 16][i % 16];
            rgb[3 * (base + i) + 2] = out.b.b[i / 16][i % 16];
        }
    }
}

/* Force the lazy cos table (dct.c:11-23) before any multi-threaded use. */
void ref_warm(void) {
    uint8_t z[64] = {0}; int o[64];
    ref_dct_block(z, o);
}
