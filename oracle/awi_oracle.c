/*
 * oracle/awi_oracle.c — TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C, strict IEEE: build with -ffp-contract=off, never
 * -ffast-math) of the BMP -> AWI encode hot path of du207/image_compression and
 * its inverse.  It is the *checker* for the CUDA path: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg may load it.  It is
 * never linked into, or called from, the product library.
 *
 * Parity status: PINNED.  The reference ships no golden vectors of its own
 * (SURVEY.md §4), so this restatement is pinned against the reference itself,
 * compiled unmodified from /root/reference/src into oracle/_ref/libawi_ref.so
 * (oracle/Makefile) — see tests/test_oracle_vs_ref.py (runs where the reference
 * build exists) and the committed vectors in tests/golden/ made by
 * oracle/make_golden.py from that build.
 *
 * Every function cites the reference file:line whose arithmetic it restates.
 * Frames are handed over the way the reference's BMP reader sees them:
 * bottom-up rows of BGR bytes, `stride` bytes apart (bmp.c:35-36,80-81,111-113).
 */
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include <math.h>
#include <stdlib.h>

#define AWI_PI 3.14159265358979323846 /* dct.h:8 */

/* ---------------------------------------------------------------- tables */

/* dct.c:86-107 — quantiser steps, [vertical freq][horizontal freq]. */
static const int QLUMA[64] = {
    16, 11, 10, 16, 24, 40, 51, 61,   12, 12, 14, 19, 26, 58, 60, 55,
    14, 13, 16, 24, 40, 57, 69, 56,   14, 17, 22, 29, 51, 87, 80, 62,
    18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55, 64, 81, 104, 113, 92,
    49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99 };
static const int QCHROMA[64] = {
    17, 18, 24, 47, 99, 99, 99, 99,   18, 21, 26, 66, 99, 99, 99, 99,
    24, 26, 56, 99, 99, 99, 99, 99,   47, 66, 99, 99, 99, 99, 99, 99,
    99, 99, 99, 99, 99, 99, 99, 99,   99, 99, 99, 99, 99, 99, 99, 99,
    99, 99, 99, 99, 99, 99, 99, 99,   99, 99, 99, 99, 99, 99, 99, 99 };

/* dct.c:147-156 — used as a GATHER index: chunk[i] = block[SCAN[i]] (dct.c:163-167). */
static const int SCAN[64] = {
     0,  1,  5,  6, 14, 15, 27, 28,    2,  4,  7, 13, 16, 26, 29, 42,
     3,  8, 12, 17, 25, 30, 41, 43,    9, 11, 18, 24, 31, 40, 44, 53,
    10, 19, 23, 32, 39, 45, 52, 54,   20, 22, 33, 38, 46, 51, 55, 60,
    21, 34, 37, 47, 50, 56, 59, 61,   35, 36, 48, 49, 57, 58, 62, 63 };

static float COS_T[8][8];
static int cos_ready = 0;

/* dct.c:14-23: (float) cos((2i+1) * j * PI / 16.0f); int*double, then / (double)16. */
static void cos_init(void) {
    if (cos_ready) return;
    for (int i = 0; i < 8; i++)
        for (int j = 0; j < 8; j++)
            COS_T[i][j] = (float) cos((double) ((2 * i + 1) * j) * AWI_PI / 16.0);
    cos_ready = 1;
}

void awi_oracle_cos_table(float out[64]) {
    cos_init();
    memcpy(out, COS_T, sizeof(COS_T));
}

void awi_oracle_tables(int luma[64], int chroma[64], int scan[64]) {
    memcpy(luma, QLUMA, sizeof(QLUMA));
    memcpy(chroma, QCHROMA, sizeof(QCHROMA));
    memcpy(scan, SCAN, sizeof(SCAN));
}

/* ---------------------------------------------------------------- colour */

/* utils.c:10-17 */
static inline uint8_t clamp_trunc(double v) {
    if (v > 255) return 255;
    if (v < 0) return 0;
    return (uint8_t) v;
}

/* color.c:31-33 — fp64, left to right, uint8 operands promoted to double. */
void awi_oracle_rgb2ycc(uint8_t r, uint8_t g, uint8_t b, uint8_t* y, uint8_t* cb, uint8_t* cr) {
    *y  = clamp_trunc(0.299 * r + 0.587 * g + 0.114 * b);
    *cb = clamp_trunc(-0.1687 * r - 0.3313 * g + 0.5 * b + 128);
    *cr = clamp_trunc(0.5 * r - 0.4187 * g - 0.0813 * b + 128);
}

/* color.c:61-63 — (cb - 128) is int arithmetic before promotion. */
void awi_oracle_ycc2rgb(uint8_t y, uint8_t cb, uint8_t cr, uint8_t* r, uint8_t* g, uint8_t* b) {
    *r = clamp_trunc(y + 1.402 * (cr - 128));
    *g = clamp_trunc(y - 0.344136 * (cb - 128) - 0.714136 * (cr - 128));
    *b = clamp_trunc(y + 1.772 * (cb - 128));
}

void awi_oracle_rgb2ycc_n(const uint8_t* rgb, size_t n, uint8_t* ycc) {
    for (size_t i = 0; i < n; i++)
        awi_oracle_rgb2ycc(rgb[3*i], rgb[3*i+1], rgb[3*i+2], &ycc[3*i], &ycc[3*i+1], &ycc[3*i+2]);
}
void awi_oracle_ycc2rgb_n(const uint8_t* ycc, size_t n, uint8_t* rgb) {
    for (size_t i = 0; i < n; i++)
        awi_oracle_ycc2rgb(ycc[3*i], ycc[3*i+1], ycc[3*i+2], &rgb[3*i], &rgb[3*i+1], &rgb[3*i+2]);
}

/* ------------------------------------------------------------- transforms */

/* dct.c:26-53.  out[u][v] = (int)(0.25*cu*cv * sum), sum (fp64) over y,x of the
 * fp32 value fl32(fl32((p-128) * cos[y][u]) * cos[x][v]).  The first product does
 * not depend on v, so it is hoisted; rounding sequence and summation order per
 * (u,v) are exactly the reference's. */
void awi_oracle_fdct(const uint8_t px[64], int coef[64]) {
    cos_init();
    double acc[64];
    for (int k = 0; k < 64; k++) acc[k] = 0.0;
    for (int y = 0; y < 8; y++) {
        for (int x = 0; x < 8; x++) {
            int8_t centred = (int8_t) (px[y * 8 + x] - 128);      /* dct.c:45 */
            for (int u = 0; u < 8; u++) {
                float first = (float) centred * COS_T[y][u];      /* fp32 multiply #1 */
                for (int v = 0; v < 8; v++) {
                    float second = first * COS_T[x][v];           /* fp32 multiply #2 */
                    acc[u * 8 + v] += (double) second;            /* fp64 accumulate */
                }
            }
        }
    }
    for (int u = 0; u < 8; u++) {
        for (int v = 0; v < 8; v++) {
            double cu = (u == 0) ? 1.0 / sqrt(2.0) : 1.0;
            double cv = (v == 0) ? 1.0 / sqrt(2.0) : 1.0;
            coef[u * 8 + v] = (int) (0.25 * cu * cv * acc[u * 8 + v]);   /* dct.c:50 */
        }
    }
}

/* dct.c:109-126: round((float)c / Q) — fp32 divide, libm round on the promoted double. */
void awi_oracle_quant(const int coef[64], int chroma, int q[64]) {
    const int* tab = chroma ? QCHROMA : QLUMA;
    for (int k = 0; k < 64; k++) q[k] = (int) round((float) coef[k] / tab[k]);
}

/* dct.c:128-145 */
void awi_oracle_dequant(const int q[64], int chroma, int coef[64]) {
    const int* tab = chroma ? QCHROMA : QLUMA;
    for (int k = 0; k < 64; k++) coef[k] = q[k] * tab[k];
}

/* dct.c:159-169 (gather) and :171-181 (scatter) */
void awi_oracle_scan(const int q[64], int chunk[64]) {
    for (int i = 0; i < 64; i++) chunk[i] = q[SCAN[i]];
}
void awi_oracle_unscan(const int chunk[64], int q[64]) {
    for (int i = 0; i < 64; i++) q[SCAN[i]] = chunk[i];
}

/* awi.c:57-62 */
void awi_oracle_transform(const uint8_t px[64], int chroma, int chunk[64]) {
    int c[64], q[64];
    awi_oracle_fdct(px, c);
    awi_oracle_quant(c, chroma, q);
    awi_oracle_scan(q, chunk);
}

/* dct.c:55-81.  fp64, terms added in (u outer, v inner) order, each term
 * (((cu*cv)*coef)*cos[y][u])*cos[x][v]; then clamp(0.25*sum + 128). */
void awi_oracle_idct(const int coef[64], uint8_t px[64]) {
    cos_init();
    for (int y = 0; y < 8; y++) {
        for (int x = 0; x < 8; x++) {
            double sum = 0.0;
            for (int u = 0; u < 8; u++) {
                for (int v = 0; v < 8; v++) {
                    double cu = (u == 0) ? 1.0 / sqrt(2.0) : 1.0;
                    double cv = (v == 0) ? 1.0 / sqrt(2.0) : 1.0;
                    sum += cu * cv * coef[u * 8 + v] * COS_T[y][u] * COS_T[x][v];
                }
            }
            px[y * 8 + x] = clamp_trunc(0.25 * sum + 128);
        }
    }
}

/* awi.c:146-151 */
void awi_oracle_inverse_transform(const int chunk[64], int chroma, uint8_t px[64]) {
    int q[64], c[64];
    awi_oracle_unscan(chunk, q);
    awi_oracle_dequant(q, chroma, c);
    awi_oracle_idct(c, px);
}

/* --------------------------------------------------------------- bit I/O */

/* bitrw.c:9-43 restated as an MSB-first accumulator over a byte buffer. */
typedef struct { uint8_t* out; size_t n; uint32_t acc; int nbits; } BitW;

static inline void bw_put(BitW* w, uint32_t value, int width) {
    for (int i = width - 1; i >= 0; i--) {
        w->acc = (w->acc << 1) | ((value >> i) & 1u);
        if (++w->nbits == 8) { w->out[w->n++] = (uint8_t) w->acc; w->acc = 0; w->nbits = 0; }
    }
}
static inline void bw_align(BitW* w) {                 /* bitrw.c:31-43 */
    if (w->nbits > 0) { w->out[w->n++] = (uint8_t) (w->acc << (8 - w->nbits)); w->acc = 0; w->nbits = 0; }
}

typedef struct { const uint8_t* in; size_t n, pos; int bit; int fail; } BitR;

static inline uint32_t br_get(BitR* r, int width) {   /* bitrw.c:46-69 */
    uint32_t v = 0;
    for (int i = 0; i < width; i++) {
        if (r->pos >= r->n) { r->fail = 1; return 0; }
        v = (v << 1) | ((r->in[r->pos] >> (7 - r->bit)) & 1u);
        if (++r->bit == 8) { r->bit = 0; r->pos++; }
    }
    return v;
}
static inline void br_align(BitR* r) {                 /* bitrw.c:72-76 */
    if (r->bit > 0) { r->bit = 0; r->pos++; }
}

static inline int bit_length(int a) {                  /* rle.c:41-50 */
    unsigned m = (unsigned) abs(a); int n = 0;
    while (m) { n++; m >>= 1; }
    return n;
}

/* rle.c:52-110.  Returns the number of bytes written (3..121 for legal input). */
int awi_oracle_rle(const int chunk[64], int prev_dc, uint8_t* out) {
    BitW w = { out, 0, 0, 0 };
    int diff = chunk[0] - prev_dc;
    bw_put(&w, (uint32_t) (diff >= 0 ? diff : 4095 + diff), 12);        /* rle.c:53-57 */
    int last = 63;
    while (last > 0 && chunk[last] == 0) last--;                         /* rle.c:60-65 */
    int run = 0;
    for (int i = 1; i <= last; i++) {
        int v = chunk[i];
        if (v == 0) {
            if (run < 15) run++;
            else { bw_put(&w, 0xF0, 8); run = 0; }                       /* rle.c:70-83 */
        } else {
            int size = bit_length(v);
            int field = v > 0 ? v : (1 << size) - 1 + v;                 /* rle.c:86-92 */
            /* RLEEntry bit-fields are 4 bits wide (rle.h:23-27); value is u16 */
            bw_put(&w, (uint32_t) (((run & 15) << 4) + (size & 15)) & 0xFFu, 8);
            bw_put(&w, (uint32_t) (uint16_t) field, size & 15);
            run = 0;
        }
    }
    bw_put(&w, 0, 8);                                                    /* rle.c:106 */
    bw_align(&w);                                                        /* rle.c:107 */
    return (int) w.n;
}

/* rle.c:113-170.  Returns bytes consumed, or -1 when the stream ends early or
 * the chunk would overrun 64 coefficients (the reference has no such bound). */
int awi_oracle_unrle(const uint8_t* in, size_t n, int prev_dc, int chunk[64]) {
    BitR r = { in, n, 0, 0, 0 };
    uint32_t raw = br_get(&r, 12);
    if (r.fail) return -1;
    int diff = raw > 2047 ? (int) raw - 4095 : (int) raw;               /* rle.c:123-125 */
    int idx = 0;
    chunk[idx++] = diff + prev_dc;
    for (;;) {
        uint32_t sym = br_get(&r, 8);
        if (r.fail) return -1;
        if (sym == 0) break;                                             /* rle.c:142 */
        int run = (sym >> 4) & 15, size = sym & 15;
        uint32_t val = br_get(&r, size);
        if (r.fail) return -1;
        if (idx + run + 1 > 64) return -1;
        for (int i = 0; i < run; i++) chunk[idx++] = 0;
        if (size > 0) chunk[idx++] = (val & (1u << (size - 1))) ? (int) val : (int) val - (1 << size) + 1;
        else chunk[idx++] = 0;                                           /* rle.c:156-158 */
    }
    while (idx < 64) chunk[idx++] = 0;
    br_align(&r);
    return (int) r.pos;
}

/* ------------------------------------------------------------ frame paths */

/* bmp.c:48-119 addressing: top-down image row y lives at bottom-up row h-1-y;
 * pixels outside w x h read as (0,0,0). */
static inline void fetch_px(const uint8_t* rows, int w, int h, size_t stride, int x, int y,
                            uint8_t* r, uint8_t* g, uint8_t* b) {
    if (x >= w || y >= h) { *r = *g = *b = 0; return; }
    const uint8_t* p = rows + (size_t) (h - 1 - y) * stride + (size_t) 3 * x;
    *b = p[0]; *g = p[1]; *r = p[2];
}

/* Six sample blocks of macroblock (mx,my): Y00 Y01 Y10 Y11 Cb Cr
 * (color.c:14-36 then the top-left pick of color.c:69-86). */
static void mb_samples(const uint8_t* rows, int w, int h, size_t stride, int mx, int my, uint8_t blk[6][64]) {
    for (int py = 0; py < 16; py++) {
        for (int px = 0; px < 16; px++) {
            uint8_t r, g, b, yy, cb, cr;
            fetch_px(rows, w, h, stride, mx * 16 + px, my * 16 + py, &r, &g, &b);
            awi_oracle_rgb2ycc(r, g, b, &yy, &cb, &cr);
            blk[(py / 8) * 2 + (px / 8)][(py % 8) * 8 + (px % 8)] = yy;
            if (!(py & 1) && !(px & 1)) {
                blk[4][(py / 2) * 8 + (px / 2)] = cb;
                blk[5][(py / 2) * 8 + (px / 2)] = cr;
            }
        }
    }
}

/* awi.c:85-128 without the 16-byte header (the host writes that, awi.c:75-83).
 * Optional outputs: qcoef (int8, n_chunks*64, scan order with ABSOLUTE DCs) and
 * chunk_len (bytes per chunk).  Encodes macroblock rows [mb_row0, mb_row1); the
 * six DC predictors at the start are those the full raster scan would hold there
 * (i.e. the preceding macroblock's DCs, found by transforming it).
 * Returns payload bytes, or (size_t)-1 if `cap` is too small. */
size_t awi_oracle_encode_rows(const uint8_t* rows, int w, int h, size_t stride,
                              int mb_row0, int mb_row1,
                              uint8_t* payload, size_t cap, int8_t* qcoef, uint32_t* chunk_len) {
    int mbw = (w + 15) / 16;
    int pred[6] = { 0, 0, 0, 0, 0, 0 };
    uint8_t blk[6][64]; int chunk[64]; uint8_t tmp[160];
    if (mb_row0 > 0) {
        mb_samples(rows, w, h, stride, mbw - 1, mb_row0 - 1, blk);
        for (int k = 0; k < 6; k++) { awi_oracle_transform(blk[k], k >= 4, chunk); pred[k] = chunk[0]; }
    }
    size_t n = 0, ci = 0;
    for (int my = mb_row0; my < mb_row1; my++) {
        for (int mx = 0; mx < mbw; mx++) {
            mb_samples(rows, w, h, stride, mx, my, blk);
            for (int k = 0; k < 6; k++, ci++) {
                awi_oracle_transform(blk[k], k >= 4, chunk);
                int len = awi_oracle_rle(chunk, pred[k], tmp);
                pred[k] = chunk[0];
                if (n + (size_t) len > cap) return (size_t) -1;
                memcpy(payload + n, tmp, (size_t) len);
                n += (size_t) len;
                if (chunk_len) chunk_len[ci] = (uint32_t) len;
                if (qcoef) for (int i = 0; i < 64; i++) qcoef[ci * 64 + i] = (int8_t) chunk[i];
            }
        }
    }
    return n;
}

size_t awi_oracle_encode_bgr(const uint8_t* rows, int w, int h, size_t stride,
                             uint8_t* payload, size_t cap, int8_t* qcoef, uint32_t* chunk_len) {
    return awi_oracle_encode_rows(rows, w, h, stride, 0, (h + 15) / 16, payload, cap, qcoef, chunk_len);
}

/* awi.c:175-217 + the two sinks: rgb_topdown == 0 -> bottom-up BGR rows `stride`
 * apart (bmp.c:222-260 and the flush order of :191-220), else packed top-down
 * RGB24 with pitch 3w (bmp.c:291-320).  Macroblock rows [mb_row0, mb_row1) only, reading from byte
 * `start` of the payload with the six predictors as they stand there (`pred_in`, NULL = zeros: the
 * frame start) — so that tests can spread a large frame over host threads; `dst` is the address of
 * the FULL frame's first byte, only the band's pixels are written.  Returns the byte offset after
 * the band's last chunk, -1 on a malformed stream. */
long awi_oracle_decode_rows(const uint8_t* payload, size_t n, int w, int h, int mb_row0, int mb_row1,
                            size_t start, const int* pred_in,
                            uint8_t* dst, size_t stride, int rgb_topdown, uint32_t* chunk_off) {
    int mbw = (w + 15) / 16;
    int pred[6] = { 0, 0, 0, 0, 0, 0 };
    if (pred_in) for (int k = 0; k < 6; k++) pred[k] = pred_in[k];
    size_t pos = start, ci = 0;
    for (int my = mb_row0; my < mb_row1; my++) {
        for (int mx = 0; mx < mbw; mx++) {
            uint8_t blk[6][64]; int chunk[64];
            for (int k = 0; k < 6; k++, ci++) {
                if (chunk_off) chunk_off[ci] = (uint32_t) pos;
                if (pos > n) return -1;
                int used = awi_oracle_unrle(payload + pos, n - pos, pred[k], chunk);
                if (used < 0) return -1;
                pos += (size_t) used;
                pred[k] = chunk[0];
                awi_oracle_inverse_transform(chunk, k >= 4, blk[k]);
            }
            for (int py = 0; py < 16; py++) {
                int y = my * 16 + py;
                if (y >= h) break;
                for (int px = 0; px < 16; px++) {
                    int x = mx * 16 + px;
                    if (x >= w) break;
                    uint8_t yy = blk[(py / 8) * 2 + (px / 8)][(py % 8) * 8 + (px % 8)];
                    uint8_t cb = blk[4][(py / 2) * 8 + (px / 2)];       /* color.c:88-105 */
                    uint8_t cr = blk[5][(py / 2) * 8 + (px / 2)];
                    uint8_t r, g, b;
                    awi_oracle_ycc2rgb(yy, cb, cr, &r, &g, &b);
                    if (rgb_topdown) {
                        uint8_t* p = dst + ((size_t) w * y + x) * 3;
                        p[0] = r; p[1] = g; p[2] = b;
                    } else {
                        uint8_t* p = dst + (size_t) (h - 1 - y) * stride + (size_t) 3 * x;
                        p[0] = b; p[1] = g; p[2] = r;
                    }
                }
            }
        }
    }
    return (long) pos;
}

long awi_oracle_decode(const uint8_t* payload, size_t n, int w, int h,
                       uint8_t* dst, size_t stride, int rgb_topdown, uint32_t* chunk_off) {
    return awi_oracle_decode_rows(payload, n, w, h, 0, (h + 15) / 16, 0, NULL, dst, stride, rgb_topdown, chunk_off);
}
