/*
 * awi_cli.c — headless counterpart of the reference CLI (main.c:22-158):
 *     awi_cli convert x.bmp     ->  x.awi
 *     awi_cli revert  x.awi     ->  x.bmp
 *     awi_cli show    x.awi     ->  x.rgb (packed top-down RGB24, the texture the reference
 *                                    uploads at main.c:226; the SDL window is left out: no
 *                                    graphics stack on this machine)
 * Exit codes follow main.c:72-86: 1 I/O or codec error, 2 wrong file suffix.
 * The BMP header handling restates bmp.c:327-343 (only 'BM' is checked, bfOffBits honoured)
 * and bmp.c:155-181 (54-byte header, 2835 px/m).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "awi_host.h"

static bool file_write(BitsOutput* self, const void* data, size_t size) {
    return fwrite(data, size, 1, (FILE*) self->context) == 1;            /* awi.c:25-33 */
}

static int ends_with(const char* s, const char* e) {
    size_t a = strlen(s), b = strlen(e);
    return a >= b && strcmp(s + a - b, e) == 0;
}

static char* swap_suffix(const char* s, const char* e) {
    size_t a = strlen(s), b = strlen(e);
    char* o = (char*) malloc(a + 1);
    strcpy(o, s);
    strcpy(o + a - b, e);
    return o;
}

static uint8_t* slurp(const char* path, size_t* n) {
    FILE* f = fopen(path, "rb");
    if (!f) return NULL;
    fseek(f, 0, SEEK_END);
    long len = ftell(f);
    fseek(f, 0, SEEK_SET);
    uint8_t* b = (uint8_t*) malloc(len > 0 ? (size_t) len : 1);
    if (b && fread(b, 1, (size_t) len, f) != (size_t) len) { free(b); b = NULL; }
    fclose(f);
    *n = (size_t) len;
    return b;
}

static int convert(const char* path) {
    if (!ends_with(path, ".bmp")) return -2;
    size_t n;
    uint8_t* bmp = slurp(path, &n);
    if (!bmp || n < 54 || bmp[0] != 'B' || bmp[1] != 'M') { fprintf(stderr, "Invalid BMP format!\n"); free(bmp); return -1; }
    uint32_t off; int32_t w, h;
    memcpy(&off, bmp + 10, 4); memcpy(&w, bmp + 18, 4); memcpy(&h, bmp + 22, 4);
    int stride = ((24 * w + 31) / 32) * 4;
    if (w <= 0 || h <= 0 || (size_t) off + (size_t) stride * h > n) { fprintf(stderr, "Reading RGB blocks error!\n"); free(bmp); return -1; }
    char* out = swap_suffix(path, ".awi");
    FILE* f = fopen(out, "wb");
    if (!f) { free(bmp); free(out); return -1; }
    BitsOutput bo = { f, 0, 0, file_write };
    bool ok = compress_rows_to_awi(bmp + off, w, h, stride, &bo);
    fclose(f);
    if (ok) printf("%s: convert completed!\n", out);
    free(out); free(bmp);
    return ok ? 0 : -1;
}

static int revert(const char* path, int rgb24) {
    if (!ends_with(path, ".awi")) return -2;
    size_t n;
    uint8_t* awi = slurp(path, &n);
    if (!awi) return -1;
    int w, h, stride;
    uint8_t* px = NULL;
    bool ok = decompress_awi_to_rows(awi, n, &w, &h, &px, &stride, rgb24);
    free(awi);
    if (!ok) return -1;
    char* out = swap_suffix(path, rgb24 ? ".rgb" : ".bmp");
    FILE* f = fopen(out, "wb");
    if (!f) { free(px); free(out); return -1; }
    if (!rgb24) {
        uint32_t img = (uint32_t) stride * h, size = 54 + img, offb = 54, z = 0, isz = 40, ppm = 2835;
        uint16_t planes = 1, bpp = 24;
        fwrite("BM", 2, 1, f); fwrite(&size, 4, 1, f); fwrite(&z, 4, 1, f); fwrite(&offb, 4, 1, f);
        fwrite(&isz, 4, 1, f); fwrite(&w, 4, 1, f); fwrite(&h, 4, 1, f); fwrite(&planes, 2, 1, f); fwrite(&bpp, 2, 1, f);
        fwrite(&z, 4, 1, f); fwrite(&img, 4, 1, f); fwrite(&ppm, 4, 1, f); fwrite(&ppm, 4, 1, f); fwrite(&z, 4, 1, f); fwrite(&z, 4, 1, f);
    }
    ok = fwrite(px, (size_t) stride * h, 1, f) == 1;
    fclose(f);
    if (ok) printf("%s: %s completed! (%dx%d)\n", out, rgb24 ? "show (headless RGB24 buffer)" : "revert", w, h);
    free(out); free(px);
    return ok ? 0 : -1;
}

int main(int argc, char** argv) {
    if (argc < 3) { fprintf(stderr, "usage: %s convert x.bmp | revert x.awi | show x.awi\n", argv[0]); return 1; }
    int rc;
    if (!strcmp(argv[1], "convert")) rc = convert(argv[2]);
    else if (!strcmp(argv[1], "revert")) rc = revert(argv[2], 0);
    else if (!strcmp(argv[1], "show")) rc = revert(argv[2], 1);
    else { fprintf(stderr, "Wrong command!\n"); return 1; }
    if (rc == -1) { fprintf(stderr, "File reading/writing error!\n"); return 1; }
    if (rc == -2) { fprintf(stderr, "Wrong file format!\n"); return 2; }
    return 0;
}
