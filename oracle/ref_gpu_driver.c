/*
 * oracle/ref_gpu_driver.c — TEST INFRASTRUCTURE: the drop-in proved at link level.
 *
 * Built only where /root/reference exists (oracle/Makefile target _ref/ref_gpu_cli): the
 * reference's OWN bmp.c, bitrw.c and utils.c and its OWN headers, with this repository's
 * host/awi_gpu.c standing in for the reference's awi.c.  The driver makes the calls of
 * main.c:104-116 (convert) and main.c:137-149 (revert): the reference's BMP reader feeds
 * 16x16 blocks to our compress_to_awi through its PixelInput vtable, our decompress_awi
 * feeds blocks to the reference's BMP writer.  main.c itself needs SDL2 and cannot be built.
 */
#include <stdio.h>
#include <string.h>

#include "awi.h"
#include "bmp.h"

/* awi.c:25-53 (the file-backed BitsOutput/BitsInput live in awi.c, which is what is being replaced) */
static bool file_write(BitsOutput* self, const void* data, size_t size) { return fwrite(data, size, 1, (FILE*) self->context) == 1; }
static bool file_read(BitsInput* self, void* buf, size_t size) { return fread(buf, size, 1, (FILE*) self->context) == 1; }

int main(int argc, char** argv) {
    if (argc != 4) { fprintf(stderr, "usage: %s convert in.bmp out.awi | revert in.awi out.bmp\n", argv[0]); return 2; }
    FILE* in = fopen(argv[2], "rb");
    FILE* out = fopen(argv[3], "wb");
    if (!in || !out) { fprintf(stderr, "File reading/writing error!\n"); return 1; }
    bool ok;
    if (!strcmp(argv[1], "convert")) {
        PixelInput pi; BMPInputContext ctx = { .fp = in };
        init_bmp_pixel_input(&pi, &ctx);
        BitsOutput bo = { .context = out, .buffer = 0, .bit_count = 0, .write = file_write };
        ok = compress_to_awi(&pi, &bo);
    } else {
        BitsInput bi = { .context = in, .buffer = 0, .bit_count = 0, .read = file_read };
        PixelOutput po; BMPOutputContext ctx = { .fp = out };
        init_bmp_pixel_output(&po, &ctx);
        ok = decompress_awi(&bi, &po);
    }
    fclose(in); fclose(out);
    return ok ? 0 : 1;
}
