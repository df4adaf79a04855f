/*
 * awi_host.h — the reference's host-side plug-in interfaces, redeclared ABI-compatibly so
 * that host/awi_gpu.c can stand in for the reference's awi.c without including its headers
 * (which do not travel with this repository).  Field order and types follow
 *   block.h:15-23   Block16_u8, RGBBlock16
 *   pixelrw.h:9-44  PixelInput, PixelOutput
 *   bitrw.h:10-43   BitsOutput, BitsInput
 *   awi.h:27-31     AWIHeader
 * A reference build can use its own headers instead: the layouts are the same.
 */
#ifndef AWI_HOST_H
#define AWI_HOST_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef AWI_USE_REFERENCE_HEADERS
/* inside the reference tree: its own declarations (awi.h pulls in bitrw.h, pixelrw.h, block.h) */
#include "awi.h"
#else

typedef struct { uint8_t b[16][16]; } Block16_u8;
typedef struct { Block16_u8 r, g, b; } RGBBlock16;

typedef struct PixelInput {
    void* context;
    int width, height;
    bool (*begin)(struct PixelInput* self);                                   /* optional */
    bool (*read_block)(struct PixelInput* self, int bx, int by, RGBBlock16* rgb_out);
    bool (*end)(struct PixelInput* self);                                     /* optional */
} PixelInput;

typedef struct PixelOutput {
    void* context;
    int width, height;
    bool (*begin)(struct PixelOutput* self, int width, int height);           /* optional */
    bool (*write_block)(struct PixelOutput* self, int bx, int by, RGBBlock16* block);
    bool (*end)(struct PixelOutput* self);                                    /* optional */
} PixelOutput;

typedef struct BitsOutput {
    void* context;
    uint8_t buffer;
    int bit_count;
    bool (*write)(struct BitsOutput* self, const void* data, size_t size);
} BitsOutput;

typedef struct BitsInput {
    void* context;
    uint8_t buffer;
    int bit_count;
    bool (*read)(struct BitsInput* self, void* buf, size_t size);
} BitsInput;

typedef struct { char title[8]; uint32_t width, height; } AWIHeader;

/* Same names, arguments and error behaviour as awi.h:45-46. */
bool compress_to_awi(PixelInput* pi, BitsOutput* bo);
bool decompress_awi(BitsInput* bi, PixelOutput* po);

#endif /* AWI_USE_REFERENCE_HEADERS */

/* Bulk variants for sources/sinks that hold the whole frame (what bmp.c's contexts can
 * provide with one fread/fwrite): no per-macroblock callbacks. */
bool compress_rows_to_awi(const uint8_t* bmp_rows, int width, int height, int row_stride, BitsOutput* bo);
/* payload = everything after the 16-byte header; rgb24 != 0 selects the packed top-down RGB sink */
bool decompress_awi_to_rows(const uint8_t* awi, size_t awi_len, int* width, int* height,
                            uint8_t** pixels, int* row_stride, int rgb24);

#endif
