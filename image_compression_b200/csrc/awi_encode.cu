// awi_encode.cu — fused BMP-rows -> AWI payload encoder for sm_100a.
//
// One kernel replaces the body of the reference's macroblock loop
// (awi.c:94-128: read_block -> rgb_block16_to_ycbcr -> ycbcr_block_subsample ->
// 6 x [dct_block -> quantize_block -> zigzag_block -> rle_encode_chunk_and_write])
// and the bit-serial writer underneath it (bitrw.c:9-43):
//
//   tile = 32 consecutive macroblocks (raster order), one CTA of 6 warps
//   A  the tile's 16 pixel rows (bottom-up BGR) -> smem: 16 bulk asynchronous copies (TMA,
//      mbarrier completion) for full interior tiles, coalesced 128-bit loads with edge
//      handling otherwise
//   B1 colour: one thread per 16-pixel row piece of a macroblock -> Y (and, on even rows, Cb/Cr of the even
//      pixels) as bytes in smem, laid out as "pair rows": 8 samples of one block next to 8 of its partner
//   B2 8 lanes per PAIR of 8x8 blocks (one row each, then one column each), both blocks of a pair in packed
//      f32x2 registers (FADD2/FMUL2/FFMA2): AAN row pass -> 8x8 transpose through smem -> AAN column pass ->
//      quantise (+ambiguity flags) with per-lane constants -> scan-order int8 coefficients in smem;
//      exact DC;  halo macroblock DCs.  One loop body serves luma and chroma pairs; the tile's 24
//      iterations (8 blocks each): four warps take the luma pairs, two the chroma pairs.
//   C  flagged coefficients recomputed warp-cooperatively in the reference's exact
//      operation order
//   D  one thread per chunk: run-length symbols + bit packing into a smem slot
//   E  CTA scan of chunk byte lengths, chunks compacted in smem
//   F  two-pass mode (regular): the tile goes to a staging buffer in completion order;
//      enc_scan_tiles + enc_gather_tiles place the tiles afterwards.
//      single-pass mode (no-alloc entry point): warp 0 resolves the tile's payload offset
//      by decoupled look-back (128 descriptors per round), coalesced word copy to it.
//
// Chunks are byte aligned (rle.c:107) so the device-wide scan is over BYTE lengths.
// The only cross-tile state, the six DC predictors (awi.c:91-92), is rebuilt from a
// one-macroblock halo with the integer DC formula, so tiles (and stripes on other
// GPUs) never wait on each other for data — and in two-pass mode not for the output offset either.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include "codec_math.cuh"
#include "awi_kernels.h"
#include "colour_fix_tables.h"
#include <mutex>
#include <type_traits>

namespace awi {

#ifndef ENC_MIN_CTAS
#define ENC_MIN_CTAS 4                           // resident CTAs per SM the register budget is sized for
#endif
constexpr int TILE_MB = ENC_TILE_MB;             // 32 macroblocks per tile
constexpr int THREADS = TILE_MB * 6;              // = chunks per tile; phase B: 4 blocks at a time per warp, 8 lanes each
constexpr int NWARPS = THREADS / 32;
constexpr int PITCH = TILE_MB * 48 + 16;         // smem bytes per pixel row: 388 words = 4 mod 32 -> the 8 rows of a
                                                 // block land 4 banks apart, conflict-free 64-bit row reads
constexpr int RAW_BYTES = 16 * PITCH;            // 24832
constexpr int SLOT = 124;                        // staging bytes per chunk (max 121), 31 words: conflict-free
constexpr int COEF_BYTES = THREADS * 64;         // int8[64] per chunk
constexpr int SAMP_MB = 400;                     // sample bytes per macroblock: 16 luma pair rows + 8 chroma pair rows of 16 B, + 16 B so
                                                 // that macroblocks sit 4 banks apart (eight of them are written side by side)
constexpr int SAMP_BYTES = TILE_MB * SAMP_MB;    // 12800
constexpr int COMPACT_CAP = (RAW_BYTES - 16) & ~15;   // the compacted tile lives in the (dead) pixel/coefficient buffer: always fits
constexpr int TR_ROW = 80;                        // phase B2 transposes, per lane group: 8 rows of 8 (block A, block B) float pairs = 64 B,
constexpr int TR_GROUP = 8 * TR_ROW;              // rows 80 B apart (20 banks: the eight 128-bit row stores of a group cover all 32 banks),
constexpr int TR_BYTES = 4 * TR_GROUP + 2 * 64;   // groups 640 B apart, 64 B more in front of groups 1, 2 (carried on) and 3 (the two groups of a half-warp
                                                  // read 64-byte rows 16 banks apart)
static_assert(THREADS == 2 * TILE_MB * 3, "loader: one fixed (macroblock, 16-byte piece) per thread, two rows per pass");
static_assert(THREADS == TILE_MB * 6, "one thread per chunk in phases D-F");
static_assert(TILE_MB % 8 == 0 && (TILE_MB * 24) % THREADS == 0, "phase B1: 24 row pieces per macroblock, whole passes");

struct __align__(16) EncSmem {
    uint8_t raw[RAW_BYTES];                      // phase A-B1: pixels; B2-D: coefficients (int8[64] per chunk, scan order); E-F: compacted tile
    uint8_t work[SAMP_BYTES + NWARPS * TR_BYTES];   // B1-C: samples [0, SAMP_BYTES), B2: transposes behind them; D-F: chunk slots
    uint8_t halo[16 * 48];                       // macroblock preceding the tile
    unsigned long long amb[THREADS];             // per chunk: natural positions to recompute
    short dc[(TILE_MB + 1) * 6];                 // block sample sums (the DCs before quantisation), [0] = halo macroblock
    int halo_sum[6];
    uint32_t warp_tot[NWARPS];
    uint8_t amb_list[THREADS];                   // chunks with flagged coefficients
    int amb_count;
    unsigned long long tile_base;
    unsigned long long load_bar;                 // mbarrier the bulk copies of phase A complete on
    int tile;
};
static_assert(THREADS * SLOT <= SAMP_BYTES + NWARPS * TR_BYTES, "chunk slots alias the sample and transpose buffers");
static_assert(COEF_BYTES <= RAW_BYTES && THREADS <= 256, "coefficients alias the pixel buffer; chunk numbers fit a byte");
static_assert(sizeof(EncSmem) + 1024 <= (228 * 1024) / ENC_MIN_CTAS, "ENC_MIN_CTAS tiles (+1 KB each the system keeps) must fit an SM's 228 KB");

// chunk index inside the tile: 6*macroblock + type, type 0..3 = Y[i][j], 4 = Cb, 5 = Cr
__device__ __forceinline__ int coef_word(int chunk, int word) { return chunk * 16 + word; }
__device__ __forceinline__ int coef_byte(int chunk, int idx) { return chunk * 64 + idx; }

// Look-back descriptors carry value and state in one 64-bit word, so relaxed accesses are
// enough (no other data is published through them) — and unlike ld.acquire they do not
// invalidate the SM's L1 on every poll.
__device__ __forceinline__ unsigned long long ld_desc(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_desc(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ---------------------------------------------------------------- phase A
// pixels are read exactly once: stream them past L1
__device__ __forceinline__ uint4 ld_stream(const uint4* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
// Bulk asynchronous copies (the TMA unit, no tensor map needed for a contiguous run): one elected lane per
// image row asks for 1536 bytes global -> shared; the bytes are counted on an mbarrier.
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, uint32_t arrivals) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(arrivals) : "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");       // the copy engine must see the initialised barrier
}
__device__ __forceinline__ void mbar_arrive_expect(unsigned long long* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
// The pixels are read exactly once: they are marked evict-first in L2, so that streaming 805 MB of them through
// does not push out the staged payload the gather pass is about to read back.
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, unsigned long long* bar) {
    unsigned long long policy;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(smem_addr(dst)), "l"(src), "r"(bytes), "r"(smem_addr(bar)), "l"(policy) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(done) : "r"(smem_addr(bar)), "r"(parity) : "memory");
    } while (!done);
}

// bmp.c:48-119 addressing.  Pixel bytes outside the image read as zero (bmp.c:89-93,106-109).
// `src` points at the piece; `valid` = bytes of it inside the image row (<= 0: none).
__device__ __forceinline__ uint4 load_piece(const uint8_t* src, int valid, int align) {
    uint4 v = make_uint4(0, 0, 0, 0);
    if (valid <= 0) return v;
    if (valid >= 16) {
        if (align == 16) return __ldg(reinterpret_cast<const uint4*>(src));
        if (align == 4) {
            const uint32_t* s = reinterpret_cast<const uint32_t*>(src);
            return make_uint4(__ldg(s), __ldg(s + 1), __ldg(s + 2), __ldg(s + 3));
        }
        valid = 16;
    }
    uint32_t w0 = 0, w1 = 0, w2 = 0, w3 = 0;
    for (int i = 0; i < valid; i++) {
        uint32_t b = (uint32_t) __ldg(src + i) << (8 * (i & 3));
        if (i < 4) w0 |= b; else if (i < 8) w1 |= b; else if (i < 12) w2 |= b; else w3 |= b;
    }
    return make_uint4(w0, w1, w2, w3);
}

// ---------------------------------------------------------------- phase B helpers
__device__ __forceinline__ int byte_of(uint32_t w, int i) { return (int) ((w >> (8 * i)) & 0xffu); }

// Where the rational quotient is not the reference's result (color.c:31-33 in fp64 lands just under an integer that
// the exact value reaches: only when the numerator is an exact multiple of the denominator) one bit per (r, g) says so,
// see colour_fix_tables.h.  A load from an 8 KB table instead of a detour through the fp64 pipe, which on this part
// kept a warp busy for hundreds of cycles while its CTA waited at the next barrier (one sample in a thousand is a
// multiple, so two warps in five met one in every pass).  Cb never needs a correction.
static __device__ const uint32_t d_y_fix[2048] = AWI_Y_FIX_BITS;
static __device__ const uint32_t d_cr_fix[2048] = AWI_CR_FIX_BITS;
__device__ __forceinline__ uint32_t fix_bit(const uint32_t* table, uint32_t i) {      // i = (r << 8) | g
    return (__ldg(table + (i >> 5)) >> (i & 31u)) & 1u;
}
// (r << 8) | g of the pixel whose B byte is byte `byte0` of the packed words w[] (a literal after unrolling):
// one PRMT where G and R share a word
__device__ __forceinline__ uint32_t rg_index(const uint32_t* w, int byte0) {
    const int wg = (byte0 + 1) >> 2, bg = (byte0 + 1) & 3, wr = (byte0 + 2) >> 2, br = (byte0 + 2) & 3;
    if (wg == wr) return __byte_perm(w[wg], 0u, (uint32_t) (bg | (br << 4) | 0x4400));
    return __byte_perm(w[wg], w[wr], (uint32_t) (bg | ((4 + br) << 4))) & 0xffffu;
}

// dp2a: c + a.lo16 * b.byte[sel] + a.hi16 * b.byte[sel+1], sel = 0 (lo) or 2 (hi); the pixel bytes
// are consumed straight from the packed words, no byte extraction.
__device__ __forceinline__ uint32_t dp2a_lo_u(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d; asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d;
}
__device__ __forceinline__ uint32_t dp2a_hi_u(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d; asm("dp2a.hi.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d;
}
__device__ __forceinline__ int dp2a_lo_s(int a, uint32_t b, int c) {      // signed 16-bit coefficients, unsigned bytes
    int d; asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d;
}
__device__ __forceinline__ int dp2a_hi_s(int a, uint32_t b, int c) {
    int d; asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d;
}
__device__ __forceinline__ int pack16(int lo, int hi) { return (lo & 0xffff) | (hi << 16); }

// luma numerator 299 r + 587 g + 114 b of the pixel whose B byte is byte `s` (0..3) of w[i]
template <int S>
__device__ __forceinline__ uint32_t luma_num(uint32_t wa, uint32_t wb) {
    constexpr uint32_t BG = 114u | (587u << 16), R0 = 299u, ZB = 114u << 16, GR = 587u | (299u << 16);
    if (S == 0) return dp2a_hi_u(R0, wa, dp2a_lo_u(BG, wa, 0u));
    if (S == 1) return dp2a_hi_u(GR, wa, dp2a_lo_u(ZB, wa, 0u));
    if (S == 2) return dp2a_lo_u(R0, wb, dp2a_hi_u(BG, wa, 0u));
    return dp2a_lo_u(GR, wb, dp2a_hi_u(ZB, wa, 0u));
}

// n / 1000 and n / 10000 for phase B1, with the exact-multiple test folded into a running minimum of the low
// product words (one 3-input minimum per two samples): an exact multiple of 1000 shows as a low word below
// 2000000; for 10000 the quotient sits in bits 35.. and only the low word is looked at (a multiple has bits
// 0..34 below 417800, so its low word is below that too; the few non-multiples that pass are told apart by
// the exact remainder test on the slow path).
__device__ __forceinline__ uint32_t div1000_q(uint32_t n, uint32_t& least_low) {
    const unsigned long long p = (unsigned long long) n * 4294968ull;
    least_low = min(least_low, (uint32_t) p);
    return (uint32_t) (p >> 32);
}
__device__ __forceinline__ uint32_t div10000_q(uint32_t n, uint32_t& least_low) {
    const unsigned long long p = (unsigned long long) n * 3435974ull;
    least_low = min(least_low, (uint32_t) p);
    return (uint32_t) (p >> 32) >> 3;
}
__device__ __forceinline__ uint32_t pack4(uint32_t a, uint32_t b, uint32_t c, uint32_t d) {      // four values 0..255 -> bytes, a lowest
    return __byte_perm(__byte_perm(a, b, 0x0040), __byte_perm(c, d, 0x0040), 0x5410);
}

// 16 pixels (48 bytes B,G,R,...) in 12 words -> their 16 luma samples (color.c:31 + utils.c:10-17) as bytes:
// .x/.y = the row of the left 8x8 block, .z/.w = of the right one
__device__ __forceinline__ uint4 luma_piece(const uint32_t* w) {
    uint32_t n[16];
#pragma unroll
    for (int q = 0; q < 4; q++) {              // 4 pixels = 3 words
        const uint32_t* x = w + 3 * q;
        n[4 * q] = luma_num<0>(x[0], x[0]); n[4 * q + 1] = luma_num<3>(x[0], x[1]);
        n[4 * q + 2] = luma_num<2>(x[1], x[2]); n[4 * q + 3] = luma_num<1>(x[2], x[2]);
    }
    // the multiple test is kept per group of four samples, so that a lane that has one looks at four samples, not sixteen
    // (measured: groups of 2 or 8 samples, one test per sample and warp votes around the slow path are all slower)
    constexpr int G = 4, NG = 16 / G;
    uint32_t f[16], least[NG];
#pragma unroll
    for (int q = 0; q < NG; q++) least[q] = 0xffffffffu;
#pragma unroll
    for (int p = 0; p < 16; p++) f[p] = div1000_q(n[p], least[p / G]);
    // Slow path, one sample in a thousand: the table lookup.
#pragma unroll
    for (int q = 0; q < NG; q++) {
        if (least[q] < 2000000u) {                     // an exact multiple of 1000 in group q (grey pixels are)
#pragma unroll
            for (int p = G * q; p < G * q + G; p++) {
                if (n[p] * 4294968u < 2000000u) f[p] -= fix_bit(d_y_fix, rg_index(w, 3 * p));
            }
        }
    }
    return make_uint4(pack4(f[0], f[1], f[2], f[3]), pack4(f[4], f[5], f[6], f[7]),
                      pack4(f[8], f[9], f[10], f[11]), pack4(f[12], f[13], f[14], f[15]));
}

// 16 pixels (48 bytes) in 12 words -> Cb (.x/.y) and Cr (.z/.w) of the 8 even pixels (color.c:32-33,82-83) as bytes
__device__ __forceinline__ uint4 chroma_piece(const uint32_t* w) {
    constexpr int BG_CB = (5000 & 0xffff) | (-3313 * 65536), R_CB = -1687 & 0xffff;
    constexpr int BG_CR = (-813 & 0xffff) | (-4187 * 65536), R_CR = 5000;
    uint32_t nb[8], nr[8];
#pragma unroll
    for (int x = 0; x < 8; x += 2) {
        const int wa = (6 * x) >> 2;               // pixels 0,4,8,12 start at byte 0 of words 0,3,6,9
        nb[x] = (uint32_t) dp2a_hi_s(R_CB, w[wa], dp2a_lo_s(BG_CB, w[wa], 1280000));
        nr[x] = (uint32_t) dp2a_hi_s(R_CR, w[wa], dp2a_lo_s(BG_CR, w[wa], 1280000));
        const int wb = (6 * (x + 1)) >> 2;         // pixels 2,6,10,14 start at byte 2 of words 1,4,7,10
        nb[x + 1] = (uint32_t) dp2a_lo_s(R_CB, w[wb + 1], dp2a_hi_s(BG_CB, w[wb], 1280000));
        nr[x + 1] = (uint32_t) dp2a_lo_s(R_CR, w[wb + 1], dp2a_hi_s(BG_CR, w[wb], 1280000));
    }
    uint32_t fb[8], fr[8], least = 0xffffffffu, unused = 0u;
#pragma unroll
    for (int x = 0; x < 8; x++) { fb[x] = div10000_q(nb[x], unused); fr[x] = div10000_q(nr[x], least); }
    if (least < 417800u) {                           // (Cb's quotient is always the reference's result: no test)
#pragma unroll
        for (int x = 0; x < 8; x++) {
            if (nr[x] - fr[x] * 10000u == 0u) fr[x] -= fix_bit(d_cr_fix, rg_index(w, 6 * x));
        }
    }
    return make_uint4(pack4(fb[0], fb[1], fb[2], fb[3]), pack4(fb[4], fb[5], fb[6], fb[7]),
                      pack4(fr[0], fr[1], fr[2], fr[3]), pack4(fr[4], fr[5], fr[6], fr[7]));
}

// single samples (halo macroblock sums)
__device__ __forceinline__ int luma_px(int r, int g, int b) {
    const uint32_t n = (uint32_t) (299 * r + 587 * g + 114 * b);
    uint32_t q = __umulhi(n, 4294968u);
    if (n - q * 1000u == 0u) q -= fix_bit(d_y_fix, ((uint32_t) r << 8) | (uint32_t) g);
    return (int) q;
}
__device__ __forceinline__ int chroma_px(int r, int g, int b, bool cr) {
    const int kr = cr ? 5000 : -1687, kg = cr ? -4187 : -3313, kb = cr ? -813 : 5000;
    const uint32_t n = (uint32_t) (kr * r + kg * g + kb * b + 1280000);
    uint32_t q = __umulhi(n, 3435974u) >> 3;
    if (cr && n - q * 10000u == 0u) q -= fix_bit(d_cr_fix, ((uint32_t) r << 8) | (uint32_t) g);
    return (int) q;
}

// Sample bytes of a tile (phase B1 -> B2, C).  Macroblock m owns SAMP_MB bytes: 16 luma pair rows (row y: 8 samples of
// the left block, 8 of the right one), then 8 chroma pair rows (row y: 8 Cb samples, 8 Cr samples).
__device__ __forceinline__ int samp_row(int mb, int type, int y) {          // byte offset of row y of the pair that holds chunk type `type`
    return mb * SAMP_MB + (type < 4 ? (((type >> 1) << 3) + y) * 16 : 256 + y * 16);
}
__device__ __forceinline__ int sample_of(const uint8_t* samp, int chunk, int y, int x) {
    const int mb = chunk / 6, type = chunk - 6 * mb;
    return samp[samp_row(mb, type, y) + ((type & 1) << 3) + x];
}
static_assert(AWI_GT_MIN_LUMA == AWI_GT_MIN_CHROMA, "phase B2 uses one ambiguity threshold for both tables");

#ifdef ENC_PHASE_CLOCKS
// debug build only (tools/phase_clocks.py): cycles per warp between phase marks, summed over all tiles.
// Kept in registers and added to one of 128 copies at the end, so the marks do not crowd one L2 slice.
__device__ unsigned long long g_phase_clk[128][16][8];
__device__ __forceinline__ long long phase_clock() { long long t; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t) :: "memory"); return t; }
#define PHASE_INIT() long long t_prev_ = phase_clock(); unsigned int ph_[13] = {}
#define PHASE_MARK(i) do { long long t_now_ = phase_clock(); ph_[i] = (unsigned int) (t_now_ - t_prev_); t_prev_ = t_now_; \
        if ((i) == 12 && lane == 0) { for (int q_ = 0; q_ < 13; q_++) atomicAdd(&g_phase_clk[blockIdx.x & 127][q_][warp], (unsigned long long) ph_[q_]); } } while (0)
extern "C" int awi_debug_phase_clocks(unsigned long long* out, int reset) {
    static unsigned long long h[128][16][8];
    if (out) {
        if (cudaMemcpyFromSymbol(h, g_phase_clk, sizeof(h)) != cudaSuccess) return 0;
        for (int i = 0; i < 128; i++) out[i] = 0;
        for (int c = 0; c < 128; c++) for (int i = 0; i < 16; i++) for (int w = 0; w < 8; w++) out[i * 8 + w] += h[c][i][w];
    }
    if (reset) { for (auto& c : h) for (auto& r : c) for (auto& v : r) v = 0; if (cudaMemcpyToSymbol(g_phase_clk, h, sizeof(h)) != cudaSuccess) return 0; }
    return 1;
}
#else
#define PHASE_MARK(i) do { } while (0)
#define PHASE_INIT() do { } while (0)
#endif

// ---------------------------------------------------------------- the kernel
__global__ void __launch_bounds__(THREADS, ENC_MIN_CTAS) awi_encode_kernel(EncParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    EncSmem& S = *reinterpret_cast<EncSmem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    pdl_wait();                       // the previous step's gather pass is done with the staging buffer and the cursor
    pdl_trigger();                    // this step's scan/gather pass may take its place on the SMs as they fall free

    // single-pass mode: ticketed tile order, so a tile only ever waits on tiles whose CTAs already run.
    // two-pass mode (staging buffer given): tiles are independent, any order will do.
    PHASE_INIT();
    const bool two_pass = P.stage != nullptr;
    int tile = (int) blockIdx.x;
    if (!two_pass) {                  // the ticket has to reach every thread: one extra barrier, single-pass mode only
        if (tid == 0) S.tile = (int) atomicAdd(P.ticket, 1u);
        __syncthreads();
        tile = S.tile;
    }
    if (tid == 0) S.amb_count = 0;    // (these are first read after the barrier that ends phase A)
    if (tid < 6) S.halo_sum[tid] = 0;
    S.amb[tid] = 0;
    // batches: tiles_per_frame tiles per frame, frames frame_stride bytes apart, predictors restart per frame
    const int frame = fast_div(tile, P.div_tpf), tile_in = tile - frame * P.tiles_per_frame;
    const int mb_first = P.mb0 + tile_in * TILE_MB;              // raster index inside the frame
    const int n_mb = min(TILE_MB, P.mb1 - mb_first);
    const bool has_halo = mb_first > 0;

    // ---- A: stage pixels.  Common case (a full tile of whole macroblocks inside the image, 16-byte aligned
    //      rows): 16 bulk copies of 1536 contiguous bytes (32 shorter ones where the tile wraps onto the next
    //      macroblock row), issued by the lanes of warp 0, land in the
    //      padded rows of S.raw on their own; nothing passes through registers.  Otherwise thread = one
    //      16-byte piece column (macroblock m, piece k) x 8 rows, with edge handling.
    {
        const uint8_t* const rows = P.rows + (long long) frame * P.frame_stride;
        const int my_first = fast_div(mb_first, P.div_mbw), mx_first = mb_first - my_first * P.mbw;
        // a full tile of whole macroblocks inside the image that spans at most two macroblock rows: every pixel row of
        // the tile is one contiguous run in the frame, or two where the tile wraps onto the next macroblock row
        const int n_a = min(TILE_MB, P.mbw - mx_first);          // macroblocks of the tile on its first macroblock row
        const bool bulk = P.align == 16 && n_mb == TILE_MB && P.mbw * 16 == P.width && TILE_MB - n_a <= P.mbw &&
                          (my_first + (n_a < TILE_MB ? 2 : 1)) * 16 <= P.height;
        if (bulk && warp == 0) {
            if (lane == 0) {
                mbar_init(&S.load_bar, 1);
                mbar_arrive_expect(&S.load_bar, 16u * TILE_MB * 48u);
            }
            __syncwarp();
            const int r = lane & 15;
            if (lane < 16) {
                const uint8_t* src = rows + (long long) (P.height - 1 - (my_first * 16 + r)) * P.stride + mx_first * 48;
                bulk_load(S.raw + r * PITCH, src, (uint32_t) n_a * 48u, &S.load_bar);
            } else if (n_a < TILE_MB) {
                const uint8_t* src = rows + (long long) (P.height - 1 - (my_first * 16 + 16 + r)) * P.stride;
                bulk_load(S.raw + r * PITCH + n_a * 48, src, (uint32_t) (TILE_MB - n_a) * 48u, &S.load_bar);
            }
        }
        const int col = tid % (3 * TILE_MB), m = col / 3, k = col - 3 * m;
        const int r0 = tid / (3 * TILE_MB);                       // 0 or 1
        uint4 v[8];
#pragma unroll
        for (int i = 0; i < 8; i++) v[i] = make_uint4(0, 0, 0, 0);
        if (!bulk && m < n_mb) {
            const int g = mb_first + m;
            const int my = fast_div(g, P.div_mbw), mx = g - my * P.mbw;
            const int xb = mx * 48 + k * 16;
            const int valid = P.width * 3 - xb;
            const int y0 = my * 16 + r0;
            const uint8_t* src = rows + (long long) (P.height - 1 - y0) * P.stride + xb;
            if (P.align == 16 && valid >= 16 && y0 + 14 < P.height) {     // 8 plain 128-bit loads
#pragma unroll
                for (int i = 0; i < 8; i++) v[i] = ld_stream(reinterpret_cast<const uint4*>(src - (long long) (2 * i) * P.stride));
            } else {
#pragma unroll
                for (int i = 0; i < 8; i++)
                    if (y0 + 2 * i < P.height) v[i] = load_piece(src - (long long) (2 * i) * P.stride, valid, P.align);
            }
        }
        // the halo macroblock's piece is requested before anything waits on the tile's own loads
        uint4 h = make_uint4(0, 0, 0, 0);
        if (tid < 48 && has_halo) {
            const int g = mb_first - 1, row = tid / 3, kk = tid - 3 * row;
            const int my = fast_div(g, P.div_mbw), mx = g - my * P.mbw;
            const int xb = mx * 48 + kk * 16, y = my * 16 + row;
            if (y < P.height)
                h = load_piece(rows + (long long) (P.height - 1 - y) * P.stride + xb, P.width * 3 - xb, P.align);
        }
        if (!bulk) {
            uint8_t* dst = S.raw + r0 * PITCH + col * 16;
#pragma unroll
            for (int i = 0; i < 8; i++) *reinterpret_cast<uint4*>(dst + i * 2 * PITCH) = v[i];
        }
        if (tid < 48) reinterpret_cast<uint4*>(S.halo)[tid] = h;
        if (bulk && warp == 0) mbar_wait(&S.load_bar, 0);          // the barrier below passes the bytes on to the other warps
    }
    PHASE_MARK(0);
    __syncthreads();
    PHASE_MARK(1);

    // ---- B1: colour.  One work item = the 16-pixel (48-byte) piece of an image row that one macroblock covers:
    //      16 items per macroblock give its luma rows, the 8 even rows once more give its (Cb, Cr) rows.  24 items per
    //      macroblock, 4 per thread.  Samples leave as bytes, 16 per item: the row of one block next to the row of its
    //      partner (left | right luma block, Cb | Cr), which is how phase B2 wants them.
    uint8_t* const samp = S.work;
    {
        constexpr int LUMA_ITEMS = TILE_MB * 16, ITEMS = TILE_MB * 24;
#pragma unroll 1
        for (int item = tid; item < ITEMS; item += THREADS) {
            if (item < LUMA_ITEMS) {                 // (warp-uniform: LUMA_ITEMS is a multiple of 32)
                const int mb = item >> 4, row = item & 15;          // eight consecutive lanes: eight rows of one macroblock, 4 banks apart
                const uint4* s4 = reinterpret_cast<const uint4*>(S.raw + row * PITCH + mb * 48);
                const uint4 a = s4[0], b = s4[1], c4 = s4[2];
                const uint32_t w[12] = { a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, c4.x, c4.y, c4.z, c4.w };
                *reinterpret_cast<uint4*>(samp + mb * SAMP_MB + row * 16) = luma_piece(w);
            } else {
                const int c = item - LUMA_ITEMS;     // eight consecutive lanes: one row of eight macroblocks (12 banks apart in, 4 out)
                const int mb = (c & 7) | ((c >> 6) << 3), crow = (c >> 3) & 7;
                const uint4* s4 = reinterpret_cast<const uint4*>(S.raw + 2 * crow * PITCH + mb * 48);
                const uint4 a = s4[0], b = s4[1], c4 = s4[2];
                const uint32_t w[12] = { a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, c4.x, c4.y, c4.z, c4.w };
                *reinterpret_cast<uint4*>(samp + mb * SAMP_MB + 256 + crow * 16) = chroma_piece(w);
            }
        }
    }
    __syncthreads();                                 // samples complete; the pixels are dead from here on

    // ---- B2: DCT + quantise.  8 lanes per PAIR of 8x8 blocks, 4 pairs (8 blocks) per warp per iteration: a lane
    //      carries row v (then column v) of two blocks in packed f32x2 registers, so every butterfly and quantiser
    //      operation is one FADD2 / FMUL2 / FFMA2 for both.  A pair's chunks are neighbours (types 0/1, 2/3, 4/5).
    //      Iterations 0..15 take the luma pairs of two macroblocks each, 16..23 the chroma pairs of four; every
    //      warp gets four of one kind.
    {
        const int v = lane & 7, jb = lane >> 3;                   // v: my row in pass 1, my column in pass 2
        uint8_t* const trg = S.work + SAMP_BYTES + warp * TR_BYTES + jb * TR_GROUP + (jb >> 1) * 64 + (jb & 1) * 64;
        uint4* const tr_row = reinterpret_cast<uint4*>(trg + v * TR_ROW);                        // my row goes out as 4 x 128 bits
        const unsigned long long* const tr_col = reinterpret_cast<const unsigned long long*>(trg + v * 8);   // my column comes back
        // per-lane quantiser constants for natural positions (u, v), u = 0..7, and where their quantised values go
        // in the chunk; reloaded when the warp moves from luma to chroma pairs
        float gk[8], gc[8];
        uint32_t pos[8];                             // shared-space address of coefficient (u, v) of chunk 0
#pragma unroll
        for (int u = 0; u < 8; u++) pos[u] = smem_addr(S.raw) + (uint32_t) unscan_rt(8 * u + v);
        constexpr int LUMA_IT = TILE_MB / 2, ALL_IT = LUMA_IT + TILE_MB / 4;
        // two warps in three take luma pairs only, the third chroma pairs only (there are twice as many luma pairs):
        // every warp loads its quantiser constants once
        constexpr int LUMA_WARPS = NWARPS * 2 / 3;
        static_assert(NWARPS % 3 == 0 && LUMA_IT % LUMA_WARPS == 0 && (ALL_IT - LUMA_IT) % (NWARPS - LUMA_WARPS) == 0 &&
                      LUMA_IT / LUMA_WARPS == (ALL_IT - LUMA_IT) / (NWARPS - LUMA_WARPS), "B2 iterations do not divide evenly");
        const bool chroma = warp >= LUMA_WARPS;
#pragma unroll
        for (int u = 0; u < 8; u++) { gk[u] = d_gk[chroma][8 * u + v]; gc[u] = d_gc[chroma][8 * u + v]; }
        const int it_end = chroma ? ALL_IT : LUMA_IT, it_step = chroma ? NWARPS - LUMA_WARPS : LUMA_WARPS;
#pragma unroll 1
        for (int it = chroma ? LUMA_IT + warp - LUMA_WARPS : warp; it < it_end; it += it_step) {
            // macroblock and pair: luma pairs are (Y00, Y01) and (Y10, Y11), chunk types 0 and 2; the chroma pair is type 4
            const int mb = chroma ? 4 * (it - LUMA_IT) + jb : 2 * it + (jb >> 1);
            const int type = chroma ? 4 : (jb & 1) << 1;
            const int chunk = mb * 6 + type;
            const bool active = mb < n_mb;
            F2 d[8];
            {
                const uint4 q = *reinterpret_cast<const uint4*>(samp + samp_row(mb, type, v));     // .x/.y block A, .z/.w block B
                // byte -> float bits 0x4B0000bb (2^23 + b) by one byte permute each; a packed subtraction of 2^23 + 128 centres both
                const F2 off = f2_splat(8388736.0f);
#pragma unroll
                for (int x = 0; x < 8; x++) {
                    const uint32_t wa = x < 4 ? q.x : q.y, wb = x < 4 ? q.z : q.w;
                    const uint32_t sel = 0x7440u + (uint32_t) (x & 3);               // 0x4B, 0x00, 0x00, byte x&3 of the word
                    d[x] = op_sub(f2_pack(__uint_as_float(__byte_perm(wa, 0x4B000000u, sel)), __uint_as_float(__byte_perm(wb, 0x4B000000u, sel))), off);
                }
            }
            aan8_t<F2>(d[0], d[1], d[2], d[3], d[4], d[5], d[6], d[7]);           // row pass
            // transpose through shared memory: a row of pairs goes out as four 128-bit stores, a column comes
            // back as eight 64-bit loads (both conflict-free, see TR_ROW)
            __syncwarp();
#pragma unroll
            for (int k = 0; k < 4; k++) {
                float x0, x1, y0, y1;
                f2_unpack(d[2 * k], x0, x1);
                f2_unpack(d[2 * k + 1], y0, y1);
                tr_row[k] = make_uint4(__float_as_uint(x0), __float_as_uint(x1), __float_as_uint(y0), __float_as_uint(y1));
            }
            __syncwarp();
#pragma unroll
            for (int u = 0; u < 8; u++) d[u].v = tr_col[u * (TR_ROW / 8)];
            aan8_t<F2>(d[0], d[1], d[2], d[3], d[4], d[5], d[6], d[7]);           // column pass: d[u] = coefficients (u, v) of both blocks

            // quantise, store (scan order, int8) and keep the largest distance from a rounding boundary
            float dist = 0.0f;
#pragma unroll
            for (int u = 0; u < 8; u++) {
                F2 t;
                const F2 df = quant_pair(d[u], gk[u], gc[u], t);
                float e0, e1, t0, t1;
                f2_unpack(df, e0, e1);
                f2_unpack(t, t0, t1);
                if (u == 0) dist = v == 0 ? 0.0f : fmaxf(fabsf(e0), fabsf(e1));      // the DC (u = v = 0) is exact, never ambiguous
                else dist = fmaxf(dist, fmaxf(fabsf(e0), fabsf(e1)));
                if (active) {
                    if (u == 0 && v == 0) {
                        // lane v = 0 holds the DC position: d[0] there is the exact integer sum of the 64 centred
                        // samples (|S| <= 8192).  It is left as a sum; phase D turns sums into quantised DCs once per
                        // chunk (dct.c:50 + :123 in integers).  The magic add reads the integer without a conversion.
                        float s0, s1;
                        f2_unpack(d[0], s0, s1);
                        short* dcp = &S.dc[(mb + 1) * 6 + type];
                        dcp[0] = (short) (__float_as_int(s0 + 12582912.0f) - 0x4B400000);
                        dcp[1] = (short) (__float_as_int(s1 + 12582912.0f) - 0x4B400000);
                    } else {
                        const uint32_t q = pos[u] + (uint32_t) chunk * 64u;          // the partner's chunk is the next one
                        asm volatile("st.shared.u8 [%0], %1;\n\tst.shared.u8 [%0+64], %2;" ::"r"(q), "r"(__float_as_uint(t0)), "r"(__float_as_uint(t1)) : "memory");
                    }
                }
            }
            const bool mine = active && dist > AWI_GT_MIN_LUMA;
            if (__any_sync(0xffffffffu, mine)) {                // rare: remember which coefficients to redo exactly
                if (mine) {
                    unsigned long long m0 = 0, m1 = 0;
#pragma unroll
                    for (int u = 0; u < 8; u++) {
                        if (u > 0 || v > 0) {
                            // here, off the fast path, every position gets its own threshold (the fast path tests the
                            // running maximum against the tightest one): two to five times fewer coefficients go to phase C
                            float x0, x1, tmp;
                            f2_unpack(d[u], x0, x1);
                            const float gt = d_gt[chroma][8 * u + v];
                            if (quant_absdiff(x0, gk[u], gc[u], tmp) > gt) m0 |= 1ull << (8 * u + v);
                            if (quant_absdiff(x1, gk[u], gc[u], tmp) > gt) m1 |= 1ull << (8 * u + v);
                        }
                    }
                    // the lane that turns a chunk's mask from zero to non-zero lists the chunk (once)
                    if (m0 && atomicOr(&S.amb[chunk], m0) == 0ull) S.amb_list[atomicAdd(&S.amb_count, 1)] = (uint8_t) chunk;
                    if (m1 && atomicOr(&S.amb[chunk + 1], m1) == 0ull) S.amb_list[atomicAdd(&S.amb_count, 1)] = (uint8_t) (chunk + 1);
                }
            }
        }
    }

    PHASE_MARK(2);
    // halo macroblock: only the six block sums are needed (integer DC formula)
    if (has_halo) {
        // 192 work items in whole warps: 128 pixel pairs for the four luma sums, 64 pixels for the two chroma sums
        for (int item = tid; item < 192; item += THREADS) {
            if (item < 128) {
                int p = 2 * item, py = p >> 4, px = p & 15;
                const uint8_t* q = S.halo + py * 48 + px * 3;
                int s = luma_px(q[2], q[1], q[0]) + luma_px(q[5], q[4], q[3]);
                s += __shfl_xor_sync(0xffffffffu, s, 1);
                s += __shfl_xor_sync(0xffffffffu, s, 2);
                s += __shfl_xor_sync(0xffffffffu, s, 8);
                s += __shfl_xor_sync(0xffffffffu, s, 16);
                if ((lane & 27) == 0) atomicAdd(&S.halo_sum[((py >> 3) << 1) | (px >> 3)], s);
            } else {
                int t = item - 128, yy = t >> 3, xx = t & 7;
                const uint8_t* q = S.halo + 2 * yy * 48 + 6 * xx;
                int cb = chroma_px(q[2], q[1], q[0], false), cr = chroma_px(q[2], q[1], q[0], true);
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    cb += __shfl_xor_sync(0xffffffffu, cb, o);
                    cr += __shfl_xor_sync(0xffffffffu, cr, o);
                }
                if (lane == 0) { atomicAdd(&S.halo_sum[4], cb); atomicAdd(&S.halo_sum[5], cr); }
            }
        }
    }
    PHASE_MARK(3);
    __syncthreads();
    PHASE_MARK(4);

    // ---- C: exact recomputation of flagged coefficients (dct.c:26-53 operation order)
    const int n_amb = S.amb_count;
    for (int e = warp; e < n_amb; e += NWARPS) {
        const int b = S.amb_list[e];
        unsigned long long m = S.amb[b];
        const bool chroma = (b % 6) >= 4;
        while (m) {
            int k = __ffsll((long long) m) - 1;
            m &= m - 1;
            int u = k >> 3, v = k & 7;
            int y = lane >> 2, x0 = (lane & 3) * 2;
            float cyu = __uint_as_float(cos_bits_rt(y * 8 + u));
            // The 64 products are multiples of 2^-28 below 2^7 (codec_math.cuh, exact path), so their sum is an integer
            // number of 2^-28 below 2^41: added up as integers, 20 low bits and the rest apart, by two warp reductions
            // (REDUX) instead of five rounds of 64-bit shuffles and fp64 additions - the whole CTA is waiting for this.
            const float p0 = exact_product(sample_of(samp, b, y, x0) - 128, cyu, __uint_as_float(cos_bits_rt(x0 * 8 + v)));
            const float p1 = exact_product(sample_of(samp, b, y, x0 + 1) - 128, cyu, __uint_as_float(cos_bits_rt((x0 + 1) * 8 + v)));
            const long long mine = __float2ll_rz(p0 * 268435456.0f) + __float2ll_rz(p1 * 268435456.0f);      // exact: powers of two
            const int lo_sum = __reduce_add_sync(0xffffffffu, (int) (mine & 0xfffff));
            const int hi_sum = __reduce_add_sync(0xffffffffu, (int) (mine >> 20));
            const double part = (double) (((long long) hi_sum << 20) + lo_sum) * 3.7252902984619140625e-9;     // * 2^-28, exact
            if (lane == 0) {
                int c = exact_coef_finish(part, u, v);
                int q = quant_int(c, chroma ? qchroma_rt(k) : qluma_rt(k));
                reinterpret_cast<int8_t*>(S.raw)[coef_byte(b, unscan_rt(k))] = (int8_t) q;
            }
        }
    }
    PHASE_MARK(5);
    if (n_amb > 0) __syncthreads();                    // (uniform) the fix-ups land before phase D; without any, the
                                                       // barrier after phase B already separates the raw pixels from the slots

    PHASE_MARK(6);
    // ---- D: run-length + bit packing, thread = chunk, into its slot
    uint32_t my_len = 0;
    const int my_mb = tid / 6, my_type = tid - 6 * my_mb;
    if (my_mb < n_mb) {
        uint32_t w[16];
#pragma unroll
        for (int q4 = 0; q4 < 4; q4++) {
            uint4 v = *reinterpret_cast<const uint4*>(S.raw + coef_byte(tid, 16 * q4));
            w[4 * q4] = v.x; w[4 * q4 + 1] = v.y; w[4 * q4 + 2] = v.z; w[4 * q4 + 3] = v.w;
        }
        uint32_t lo = 0, hi = 0;
#pragma unroll
        for (int i = 15; i >= 0; i--) {
            // 0x80 in every non-zero byte; one multiply gathers the four flags (bits 7, 15, 23, 31) in bits 28..31 (the
            // partial products below bit 24 cannot carry that far); a funnel shift pushes the nibble into the mask
            const uint32_t nz = (((w[i] & 0x7f7f7f7fu) + 0x7f7f7f7fu) | w[i]) & 0x80808080u;
            const uint32_t g = nz * 0x00204081u;
            if (i < 8) lo = __funnelshift_l(g, lo, 4); else hi = __funnelshift_l(g, hi, 4);
        }
        const int8_t* cbase = reinterpret_cast<const int8_t*>(S.raw);
        auto coef = [&](int i) -> int { return (int) cbase[coef_byte(tid, i)]; };
        uint32_t* slot = reinterpret_cast<uint32_t*>(S.work + tid * SLOT);
        auto put = [&](int word, uint32_t be) { slot[word] = __byte_perm(be, 0, 0x0123); };
        const int c_dc = dc_coef_from_sum(S.dc[(my_mb + 1) * 6 + my_type]), 
                  c_prev = dc_coef_from_sum(my_mb > 0 ? S.dc[my_mb * 6 + my_type]                  // previous macroblock of the tile
                                                      : (has_halo ? S.halo_sum[my_type] - 128 * 64 : 0));   // or the halo macroblock / frame start
        const int dc = my_type >= 4 ? quant_int(c_dc, 17) : quant_int(c_dc, 16);             // constant divisors
        const int prev = my_type >= 4 ? quant_int(c_prev, 17) : quant_int(c_prev, 16);
        reinterpret_cast<int8_t*>(S.raw)[coef_byte(tid, 0)] = (int8_t) dc;       // for the optional coefficient output
        my_len = (uint32_t) rle_chunk(dc, prev, ((unsigned long long) hi << 32) | lo, coef, put);
    }

    PHASE_MARK(7);
    // ---- E: offsets inside the tile
    uint32_t total, my_off;
    {
        uint32_t incl = my_len;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) S.warp_tot[warp] = incl;
        PHASE_MARK(8);
        __syncthreads();                                // also: every slot written, coefficients no longer read by D
        uint32_t wbase = 0;
        total = 0;
#pragma unroll
        for (int i = 0; i < NWARPS; i++) { uint32_t t = S.warp_tot[i]; if (i < warp) wbase += t; total += t; }
        my_off = wbase + incl - my_len;
    }
    PHASE_MARK(9);
    // optional per-chunk outputs (tests, decode side-info) need the coefficient buffer: before it is reused
    if (P.qcoef != nullptr) {
        if (my_mb < n_mb) {
            uint32_t* dst = reinterpret_cast<uint32_t*>(P.qcoef + ((long long) frame * (P.mb1 - P.mb0) * 6 + (long long) (mb_first - P.mb0) * 6 + tid) * 64);
#pragma unroll
            for (int i = 0; i < 16; i++) dst[i] = reinterpret_cast<const uint32_t*>(S.raw)[coef_word(tid, i)];
        }
        __syncthreads();
    }

    const bool compact_ok = total + 16 <= (uint32_t) P.compact_cap;   // normally COMPACT_CAP; smaller only under the self-test hook
    uint8_t* compact = S.raw;

    if (two_pass) {
        // ---- two-pass tail: no tile waits for another.  Reserve room in the staging buffer (16-byte units,
        // completion order), leave (where, how many bytes) for enc_scan_tiles / enc_gather_tiles.
        if (tid == 0) {
            const unsigned units = (total + 15u) >> 4;
            // a staging buffer that holds the worst case of every tile gives each tile its own slot: no round trip to
            // the cursor in global memory (about a microsecond during which the CTA's other threads wait at the barrier below)
            const unsigned at = P.stage_slot_units ? (unsigned) tile * P.stage_slot_units : atomicAdd(P.stage_cursor, units);
            S.tile_base = at;
            P.tile_info[tile] = make_uint2(at, total);
            if (P.bucket != nullptr) atomicAdd(P.bucket + (tile >> 6), total);      // the gather pass adds up 64 tiles at a time
            if ((unsigned long long) at + units > P.stage_cap_units) atomicOr(P.err, 1u);
        }
        if (compact_ok) {
            const uint32_t* src = reinterpret_cast<const uint32_t*>(S.work + tid * SLOT);      // (slots are word aligned)
            uint8_t* dst = compact + my_off;
#pragma unroll 1
            for (uint32_t i = 0; i < my_len; i += 4) {
                const uint32_t w = src[i >> 2];
                dst[i] = (uint8_t) w;
                if (i + 1 < my_len) dst[i + 1] = (uint8_t) (w >> 8);
                if (i + 2 < my_len) dst[i + 2] = (uint8_t) (w >> 16);
                if (i + 3 < my_len) dst[i + 3] = (uint8_t) (w >> 24);
            }
        }
        PHASE_MARK(10);
        __syncthreads();
        PHASE_MARK(11);
        const unsigned long long at = S.tile_base;
        const bool fits = at + ((total + 15u) >> 4) <= P.stage_cap_units;
        if (P.chunk_off != nullptr && my_mb < n_mb)
            P.chunk_off[(long long) frame * (P.mb1 - P.mb0) * 6 + (long long) (mb_first - P.mb0) * 6 + tid] = my_off;          // tile-relative; the gather pass adds the tile's offset
        if (fits) {
            uint8_t* out = P.stage + at * 16;
            if (compact_ok) {
                const uint4* c4 = reinterpret_cast<const uint4*>(compact);
                uint4* o4 = reinterpret_cast<uint4*>(out);
#pragma unroll 1
                for (uint32_t j = tid; j < ((total + 15u) >> 4); j += THREADS) o4[j] = c4[j];
            } else {                                     // oversized tile (> 64 B per chunk on average): chunk by chunk
                const uint8_t* src = S.work + tid * SLOT;
#pragma unroll 1
                for (uint32_t i = 0; i < my_len; i++) out[my_off + i] = src[i];
            }
        }
        PHASE_MARK(12);
        return;
    }

    // ---- single-pass tail.  warp 0: decoupled look-back for the tile's payload offset, 4 descriptors per lane per round
    if (warp == 0) {
        constexpr unsigned long long FLAG_AGG = 1ull << 62, FLAG_PFX = 2ull << 62, VMASK = (1ull << 62) - 1;
        unsigned long long excl = 0;
        if (tile > 0) {
            if (lane == 0) st_desc(P.desc + tile, FLAG_AGG | total);
            int base = tile - 1;
            while (true) {
                unsigned long long dsc[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int idx = base - 4 * lane - j;
                    dsc[j] = idx >= 0 ? ld_desc(P.desc + idx) : FLAG_PFX;      // virtual prefix 0 before tile 0
                }
                // only descriptors up to the nearest prefix matter; wait for those
                unsigned long long sum = 0;
                bool found = false;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    if (!found) {
                        const int idx = base - 4 * lane - j;
                        while ((dsc[j] >> 62) == 0) { __nanosleep(40); dsc[j] = ld_desc(P.desc + idx); }
                        sum += dsc[j] & VMASK;
                        found = (dsc[j] >> 62) == 2;
                    }
                }
                const unsigned pm = __ballot_sync(0xffffffffu, found);
                const int first = pm ? (__ffs((int) pm) - 1) : 31;
                unsigned long long c = lane <= first ? sum : 0ull;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
                excl += c;
                if (pm) break;
                base -= 128;
            }
        }
        if (lane == 0) {
            st_desc(P.desc + tile, FLAG_PFX | (excl + total));
            S.tile_base = excl;
            if (tile == P.n_tiles - 1) *P.out_len = excl + total;         // last tile: payload length
        }
    }
    // meanwhile / afterwards: every thread moves its own chunk to its place in the compacted tile
    if (compact_ok) {
        const uint8_t* src = S.work + tid * SLOT;
        uint8_t* dst = compact + my_off;
#pragma unroll 1
        for (uint32_t i = 0; i < my_len; i++) dst[i] = src[i];
    }
    __syncthreads();

    // ---- F: the tile's bytes go to their final place
    {
        const unsigned long long base = S.tile_base;
        const bool fits = base + total <= P.out_cap;
        if (!fits && tid == 0) atomicOr(P.err, 1u);
        if (P.chunk_off != nullptr && my_mb < n_mb)
            P.chunk_off[(long long) frame * (P.mb1 - P.mb0) * 6 + (long long) (mb_first - P.mb0) * 6 + tid] = (uint32_t) (base + my_off);
        if (fits && compact_ok) {
            uint8_t* out = P.out + base;
            const uint32_t head = min(total, (uint32_t) ((4 - (reinterpret_cast<uintptr_t>(out) & 3)) & 3));
            if (tid < head) out[tid] = compact[tid];
            const uint32_t n_words = (total - head) >> 2;
            const uint32_t* cw = reinterpret_cast<const uint32_t*>(compact);
            uint32_t* ow = reinterpret_cast<uint32_t*>(out + head);
            const uint32_t sh = head * 8;
#pragma unroll 1
            for (uint32_t j = tid; j < n_words; j += THREADS) {
                // bytes compact[head + 4j .. +4): two aligned words, funnel-shifted
                uint32_t lo = cw[j], hi = cw[j + 1];
                ow[j] = __funnelshift_r(lo, hi, sh);
            }
            const uint32_t tail0 = head + (n_words << 2);
            if (tail0 + tid < total) out[tail0 + tid] = compact[tail0 + tid];
        } else if (fits) {                              // oversized tile: chunk by chunk
            const uint8_t* src = S.work + tid * SLOT;
            uint8_t* dst = P.out + base + my_off;
#pragma unroll 1
            for (uint32_t i = 0; i < my_len; i++) dst[i] = src[i];
        }
    }
}

// ---------------------------------------------------------------- two-pass mode, passes 2 and 3
// Exclusive scan of the tiles' byte counts in tile order (one CTA of 32 warps; each warp owns a
// contiguous range and reads it 32 tiles at a time, coalesced).
__global__ void __launch_bounds__(1024) enc_scan_tiles(const uint2* __restrict__ tile_info, int n_tiles,
                                                       unsigned long long* __restrict__ tile_off,
                                                       unsigned long long* __restrict__ out_len,
                                                       const unsigned long long* __restrict__ chain_in,
                                                       unsigned long long* __restrict__ chain_out,
                                                       unsigned long long out_cap, unsigned int* err,
                                                       unsigned long long* __restrict__ out_len_user,
                                                       unsigned int* __restrict__ stage_cursor) {
    __shared__ unsigned long long wsum[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    pdl_wait();
    pdl_trigger();
    const int per = ((n_tiles + 31) / 32 + 31) & ~31;               // tiles per warp, a multiple of 32
    const int t0 = min(n_tiles, warp * per), t1 = min(n_tiles, t0 + per);
    unsigned long long mine = 0;
    for (int t = t0 + lane; t < t1; t += 32) mine += tile_info[t].y;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
    if (lane == 0) wsum[warp] = mine;
    __syncthreads();
    if (warp == 0) {
        unsigned long long w = wsum[lane], wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned long long v = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += v;
        }
        // chained calls (sub-stripes of one frame) continue where the previous call's payload ended
        const unsigned long long base = chain_in != nullptr ? *chain_in : 0ull;
        wsum[lane] = base + wi - w;
        if (lane == 31) {
            *out_len = base + wi;
            if (out_len_user != nullptr) *out_len_user = base + wi;
            *stage_cursor = 0u;            // every tile of this launch has reserved its room: ready for the next launch
            if (chain_out != nullptr) *chain_out = base + wi;
            if (base + wi > out_cap) atomicOr(err, 1u);
        }
    }
    __syncthreads();
    unsigned long long run = wsum[warp];
    for (int i0 = t0; i0 < t1; i0 += 32) {
        const int t = i0 + lane;
        const unsigned long long v = t < t1 ? (unsigned long long) tile_info[t].y : 0ull;
        unsigned long long incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned long long u = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += u;
        }
        if (t < t1) tile_off[t] = run + incl - v;
        run += __shfl_sync(0xffffffffu, incl, 31);
    }
}

// One warp per tile: staging (16-byte aligned, completion order) -> payload (byte offsets, tile order).
__global__ void __launch_bounds__(256) enc_gather_tiles(const uint8_t* __restrict__ stage, const uint2* __restrict__ tile_info,
                                                        const unsigned long long* __restrict__ tile_off, int n_tiles,
                                                        uint8_t* __restrict__ out, const unsigned long long* __restrict__ out_len,
                                                        unsigned long long out_cap, uint32_t* __restrict__ chunk_off,
                                                        int tiles_per_frame, int chunks_per_frame,
                                                        unsigned long long* __restrict__ frame_off,
                                                        unsigned long long stage_cap_units) {
    const int tile = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    pdl_wait();
    pdl_trigger();
    if (tile >= n_tiles || *out_len > out_cap) return;
    const uint2 info = tile_info[tile];
    if ((unsigned long long) info.x + ((info.y + 15u) >> 4) > stage_cap_units) return;   // never staged (err bit 0 is already set)
    const unsigned long long off = tile_off[tile];
    const uint32_t total = info.y;
    const uint8_t* src = stage + (unsigned long long) info.x * 16;
    uint8_t* dst = out + off;
    const uint32_t head = min(total, (uint32_t) ((4 - (reinterpret_cast<uintptr_t>(dst) & 3)) & 3));
    if (lane < head) dst[lane] = src[lane];
    const uint32_t n_words = (total - head) >> 2;
    const uint32_t* sw = reinterpret_cast<const uint32_t*>(src);
    uint32_t* dw = reinterpret_cast<uint32_t*>(dst + head);
    const uint32_t sh = head * 8;
    for (uint32_t j = lane; j < n_words; j += 32) dw[j] = __funnelshift_r(sw[j], sw[j + 1], sh);
    const uint32_t tail0 = head + (n_words << 2);
    if (tail0 + lane < total) dst[tail0 + lane] = src[tail0 + lane];
    const int frame = tile / tiles_per_frame, tile_in = tile - frame * tiles_per_frame;
    if (frame_off != nullptr && lane == 0) {
        if (tile_in == 0) frame_off[frame] = off;                              // where each frame's payload starts
        if (tile == n_tiles - 1) frame_off[frame + 1] = off + total;           // ... and where the last one ends
    }
    if (chunk_off != nullptr) {
        const int c_in = tile_in * (TILE_MB * 6);
        const long long c0 = (long long) frame * chunks_per_frame + c_in;
        for (int i = lane; i < TILE_MB * 6 && c_in + i < chunks_per_frame; i += 32) chunk_off[c0 + i] += (uint32_t) off;
    }
}

// Frames of up to ENC_FUSED_MAX_TILES tiles (a 4K or 8K frame, a sub-stripe, a stripe of a frame spread over
// several GPUs): the scan is folded into the gather pass — every CTA adds up the byte counts in front of its
// eight tiles itself (at most 4096 coalesced loads, from L2), which saves a launch on frames where a launch is
// a tenth of the work.
constexpr int ENC_FUSED_MAX_TILES = 4096;
__global__ void __launch_bounds__(256) enc_scan_gather_tiles(const uint8_t* __restrict__ stage, const uint2* __restrict__ tile_info,
                                                             int n_tiles, uint8_t* __restrict__ out,
                                                             unsigned long long* __restrict__ out_len,
                                                             unsigned long long* __restrict__ out_len_user,
                                                             const unsigned long long* __restrict__ chain_in,
                                                             unsigned long long* __restrict__ chain_out,
                                                             unsigned long long out_cap, unsigned int* err,
                                                             unsigned int* __restrict__ stage_cursor,
                                                             uint32_t* __restrict__ chunk_off, int tiles_per_frame,
                                                             int chunks_per_frame, unsigned long long* __restrict__ frame_off,
                                                             unsigned long long stage_cap_units,
                                                             unsigned int* __restrict__ bucket, unsigned int* __restrict__ gather_done) {
    __shared__ unsigned long long wsum[8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    pdl_wait();                       // every tile of the encode kernel is staged and recorded
    pdl_trigger();
    const int first = blockIdx.x * 8;
    unsigned long long mine = 0;
    if (bucket != nullptr) {          // larger frames: the encode kernel has added the byte counts up per 64 tiles
        const int b = first >> 6;
        for (int t = tid; t < b; t += 256) mine += bucket[t];
        for (int t = (b << 6) + tid; t < first; t += 256) mine += tile_info[t].y;
    } else {
        for (int t = tid; t < first; t += 256) mine += tile_info[t].y;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
    if (lane == 0) wsum[warp] = mine;
    // my tile's size and the sizes of the CTA's tiles in front of it
    const int tile = first + warp;
    const uint32_t sz = (lane < 8 && first + lane < n_tiles) ? tile_info[first + lane].y : 0u;
    uint32_t before = lane < warp ? sz : 0u;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) before += __shfl_xor_sync(0xffffffffu, before, o);
    before = __shfl_sync(0xffffffffu, before, 0);
    const uint32_t total = __shfl_sync(0xffffffffu, sz, warp);
    __syncthreads();
    unsigned long long off = chain_in != nullptr ? *chain_in : 0ull;
#pragma unroll
    for (int w = 0; w < 8; w++) off += wsum[w];
    off += before;
    if (bucket != nullptr && tid == 0) {                      // the CTA that reads the sums last leaves them at zero for the next launch
        __threadfence();
        if (atomicAdd(gather_done, 1u) == gridDim.x - 1) {
            for (int t = 0; t < (n_tiles + 63) >> 6; t++) bucket[t] = 0u;
            *gather_done = 0u;
        }
    }
    if (tile >= n_tiles) return;
    const bool fits = off + total <= out_cap;
    if (tile == n_tiles - 1 && lane == 0) {                  // the end of the last tile is the end of the payload
        *out_len = off + total;
        if (out_len_user != nullptr) *out_len_user = off + total;
        if (chain_out != nullptr) *chain_out = off + total;
        *stage_cursor = 0u;                                  // ready for the next launch
        if (!fits) atomicOr(err, 1u);
    }
    const int frame = tile / tiles_per_frame, tile_in = tile - frame * tiles_per_frame;
    if (frame_off != nullptr && lane == 0) {
        if (tile_in == 0) frame_off[frame] = off;
        if (tile == n_tiles - 1) frame_off[frame + 1] = off + total;
    }
    if (chunk_off != nullptr) {
        const int c_in = tile_in * (TILE_MB * 6);
        const long long c0 = (long long) frame * chunks_per_frame + c_in;
        for (int i = lane; i < TILE_MB * 6 && c_in + i < chunks_per_frame; i += 32) chunk_off[c0 + i] += (uint32_t) off;
    }
    if (!fits) return;                                       // (offsets only grow: the last tile reports it)
    // Tiles are staged in completion order, so a tile from the front of the frame can have reserved room
    // behind the end of the staging buffer (and then staged nothing) while its payload offset still fits:
    // it must not be copied from there.  The encode kernel has raised err bit 0 for it already.
    const unsigned long long at = tile_info[tile].x;
    if (at + ((total + 15u) >> 4) > stage_cap_units) return;
    const uint8_t* src = stage + at * 16;
    uint8_t* dst = out + off;
    const uint32_t head = min(total, (uint32_t) ((4 - (reinterpret_cast<uintptr_t>(dst) & 3)) & 3));
    if (lane < head) dst[lane] = src[lane];
    const uint32_t n_words = (total - head) >> 2;
    const uint32_t* sw = reinterpret_cast<const uint32_t*>(src);
    uint32_t* dw = reinterpret_cast<uint32_t*>(dst + head);
    const uint32_t sh = head * 8;
    for (uint32_t j = lane; j < n_words; j += 32) dw[j] = __funnelshift_r(sw[j], sw[j + 1], sh);
    const uint32_t tail0 = head + (n_words << 2);
    if (tail0 + lane < total) dst[tail0 + lane] = src[tail0 + lane];
}

// ---------------------------------------------------------------- launch
size_t enc_smem_bytes() { return sizeof(EncSmem); }

bool pdl_enabled() {
    static const bool on = getenv("GPU_CODEC_NO_PDL") == nullptr;
    return on;
}

cudaError_t launch_encode(const EncParams& P, int n_tiles, cudaStream_t stream) {
    static bool attr_set[64] = { false };
    static std::mutex attr_mu;                       // contexts of several devices may launch from different host threads
    int dev = 0;
    cudaGetDevice(&dev);
    std::unique_lock<std::mutex> attr_lock(attr_mu);
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(awi_encode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int) sizeof(EncSmem));
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(awi_encode_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                 cudaSharedmemCarveoutMaxShared);
        if (e != cudaSuccess) return e;
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    attr_lock.unlock();
    EncParams Q = P;
    Q.n_tiles = n_tiles;
    if (Q.compact_cap <= 0 || Q.compact_cap > COMPACT_CAP) Q.compact_cap = COMPACT_CAP;
    if (Q.tiles_per_frame <= 0) { Q.tiles_per_frame = n_tiles; Q.frame_stride = 0; }      // a single frame
    Q.div_mbw = make_fastdiv((uint32_t) Q.mbw);
    Q.div_tpf = make_fastdiv((uint32_t) Q.tiles_per_frame);
    if (n_tiles <= ENC_FUSED_MAX_TILES || n_tiles > 64 * ENC_BUCKETS) Q.bucket = nullptr;   // (the kernel adds into them only where the gather pass reads them)
    cudaError_t le = launch_pdl(awi_encode_kernel, (unsigned) n_tiles, THREADS, sizeof(EncSmem), stream, Q);
    if (le != cudaSuccess) return le;
    if (P.stage != nullptr) {
        const bool buckets = n_tiles > ENC_FUSED_MAX_TILES && P.bucket != nullptr && n_tiles <= 64 * ENC_BUCKETS;
        if (n_tiles <= ENC_FUSED_MAX_TILES || buckets) {
            le = launch_pdl(enc_scan_gather_tiles, (unsigned) (n_tiles + 7) / 8, 256, 0, stream, P.stage, P.tile_info, n_tiles, P.out, P.out_len, P.out_len_user,
                                                                         P.chain_in, P.chain_out, P.out_cap, P.err, P.stage_cursor,
                                                                         P.chunk_off, Q.tiles_per_frame, (P.mb1 - P.mb0) * 6, P.frame_off,
                                                                         P.stage_cap_units, buckets ? P.bucket : (unsigned int*) nullptr, P.gather_done);
        } else {
            le = launch_pdl(enc_scan_tiles, 1, 1024, 0, stream, P.tile_info, n_tiles, P.tile_off, P.out_len, P.chain_in, P.chain_out, P.out_cap, P.err,
                                                   P.out_len_user, P.stage_cursor);
            if (le == cudaSuccess) le = launch_pdl(enc_gather_tiles, (unsigned) (n_tiles + 7) / 8, 256, 0, stream, P.stage, P.tile_info, P.tile_off, n_tiles, P.out, P.out_len,
                                                                     P.out_cap, P.chunk_off, Q.tiles_per_frame,
                                                                     (P.mb1 - P.mb0) * 6, P.frame_off, P.stage_cap_units);
        }
    }
    return le != cudaSuccess ? le : cudaGetLastError();
}

}  // namespace awi
