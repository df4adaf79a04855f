// tests/emul/emul.cpp — TEST HARNESS: compiles image_compression_b200/csrc/codec_math.cuh
// for the host so the fast-path + exact-fallback arithmetic the CUDA kernels use can be
// checked against the oracle on a box without a GPU.  Not part of the product: the
// product library has no CPU path.  Build: g++ -O2 -ffp-contract=off -shared -fPIC.
#include <stdint.h>
#include <string.h>
#include "../../image_compression_b200/csrc/codec_math.cuh"

using namespace awi;

static float cosf_tab(int i, int j) { return u2f(cos_bits(i * 8 + j)); }

// one block: samples (0..255) -> chunk (scan order) the way the encode kernel does it.
// returns the number of positions that took the exact path.
template <bool CHROMA>
static int transform(const uint8_t* px, int* chunk) {
    float d[64]; uint32_t packed[16]; uint64_t amb;
    for (int i = 0; i < 64; i++) d[i] = centred(px[i]);
    fdct_quant_pack<CHROMA>(d, packed, amb);
    int8_t q[64];
    memcpy(q, packed, 64);
    int n = 0;
    for (int k = 1; k < 64; k++) {
        if (!((amb >> k) & 1)) continue;
        n++;
        int u = k >> 3, v = k & 7;
        double sum = 0.0;
        for (int y = 0; y < 8; y++)
            for (int x = 0; x < 8; x++)
                sum += (double) exact_product((int) px[y * 8 + x] - 128, cosf_tab(y, u), cosf_tab(x, v));
        int c = exact_coef_finish(sum, u, v);
        q[unscan_tab(k)] = (int8_t) quant_int(c, CHROMA ? qchroma(k) : qluma(k));
    }
    for (int i = 0; i < 64; i++) chunk[i] = q[i];
    return n;
}

extern "C" {

// Checks fast_div(n, make_fastdiv(d)) == n / d for every n in the list and every d in [d0, d1); returns the first
// failing d (and its n through *bad_n), or 0.
uint32_t emul_fastdiv_check(uint32_t d0, uint32_t d1, const int32_t* ns, int n_ns, int32_t* bad_n) {
    for (uint32_t d = d0; d < d1; d++) {
        const FastDiv f = make_fastdiv(d);
        for (int i = 0; i < n_ns; i++) {
            int32_t n = ns[i];
            if (fast_div(n, f) != (int) ((uint32_t) n / d)) { *bad_n = n; return d; }
        }
        // the neighbourhood of multiples of d near the top of the range, where an inexact multiplier fails first
        for (uint32_t q = 0x7fffffffu / d, k = 0; k < 3 && q > 0; k++, q--)
            for (int e = -1; e <= 0; e++) {
                const uint64_t n64 = (uint64_t) q * d + (uint64_t) (int64_t) e;
                if (n64 > 0x7fffffffull) continue;
                const int32_t n = (int32_t) n64;
                if (fast_div(n, f) != (int) ((uint32_t) n / d)) { *bad_n = n; return d; }
            }
    }
    return 0;
}

int emul_transform(const uint8_t* px, int chroma, int* chunk) {
    return chroma ? transform<true>(px, chunk) : transform<false>(px, chunk);
}

// n blocks; returns total exact-path count; chunks: n*64 ints
long emul_transform_n(const uint8_t* px, long n, int chroma, int* chunks) {
    long tot = 0;
    for (long b = 0; b < n; b++) tot += emul_transform(px + 64 * b, chroma, chunks + 64 * b);
    return tot;
}

// fast path only (no fallback) + ambiguity mask, for measuring how good the flagging is
void emul_fast_only(const uint8_t* px, int chroma, int* chunk, uint64_t* amb_out) {
    float d[64]; uint32_t packed[16]; uint64_t amb;
    for (int i = 0; i < 64; i++) d[i] = centred(px[i]);
    if (chroma) fdct_quant_pack<true>(d, packed, amb); else fdct_quant_pack<false>(d, packed, amb);
    int8_t q[64]; memcpy(q, packed, 64);
    for (int i = 0; i < 64; i++) chunk[i] = q[i];
    *amb_out = amb;
}

// AAN outputs of the fast path (before the quantiser; scaled by 8 a_u a_v relative to the reference's value) next to
// the reference's exact pre-truncation value ((0.25 cu cv) * sum, dct.c:50) for every position: the test compares
// their distance with the a-priori bound AWI_EPS_T the ambiguity thresholds are built from.
void emul_fast_vs_exact(const uint8_t* px, double* fast_scaled, double* exact_t) {
    float d[64];
    for (int i = 0; i < 64; i++) d[i] = centred(px[i]);
    for (int y = 0; y < 8; y++)
        aan8(d[8 * y + 0], d[8 * y + 1], d[8 * y + 2], d[8 * y + 3], d[8 * y + 4], d[8 * y + 5], d[8 * y + 6], d[8 * y + 7]);
    for (int x = 0; x < 8; x++)
        aan8(d[x], d[8 + x], d[16 + x], d[24 + x], d[32 + x], d[40 + x], d[48 + x], d[56 + x]);
    for (int k = 0; k < 64; k++) {
        int u = k >> 3, v = k & 7;
        double sum = 0.0;
        for (int y = 0; y < 8; y++)
            for (int x = 0; x < 8; x++)
                sum += (double) exact_product((int) px[y * 8 + x] - 128, cosf_tab(y, u), cosf_tab(x, v));
        fast_scaled[k] = (double) d[k];
        exact_t[k] = dmul_rn(fdct_scale(u, v), sum);
    }
}

// The encode kernel's own flag expression (awi_encode.cu phase B2), two levels: a lane (one column v of the block) keeps
// the running maximum of |g - rn(g)| over its eight positions and compares it with ONE threshold, the tightest of the
// table (AWI_GT_MIN_*); only a lane that exceeds it looks at its positions one by one, each against its own threshold
// from the table.  The DC position is never flagged.  Returns the fast-path chunk and the flag mask (natural positions).
void emul_fast_kernel_flags(const uint8_t* px, int chroma, int* chunk, uint64_t* amb_out) {
    float d[64];
    for (int i = 0; i < 64; i++) d[i] = centred(px[i]);
    for (int y = 0; y < 8; y++)
        aan8(d[8 * y + 0], d[8 * y + 1], d[8 * y + 2], d[8 * y + 3], d[8 * y + 4], d[8 * y + 5], d[8 * y + 6], d[8 * y + 7]);
    for (int x = 0; x < 8; x++)
        aan8(d[x], d[8 + x], d[16 + x], d[24 + x], d[32 + x], d[40 + x], d[48 + x], d[56 + x]);
    const float gt_min = chroma ? AWI_GT_MIN_CHROMA : AWI_GT_MIN_LUMA;
    uint64_t amb = 0;
    int q[64];
    float dist[64];
    dist[0] = 0.0f;
    for (int k = 1; k < 64; k++) {
        float t;
        dist[k] = quant_absdiff(d[k], chroma ? gk_chroma(k) : gk_luma(k), chroma ? gc_chroma(k) : gc_luma(k), t);
        q[k] = (int) (int8_t) (f2u(t) & 0xffu);
    }
    for (int v = 0; v < 8; v++) {
        float worst = 0.0f;
        for (int u = 0; u < 8; u++) worst = dist[8 * u + v] > worst ? dist[8 * u + v] : worst;
        if (worst > gt_min)
            for (int u = 0; u < 8; u++) {
                const int k = 8 * u + v;
                if (k > 0 && dist[k] > (chroma ? gt_chroma(k) : gt_luma(k))) amb |= 1ull << k;
            }
    }
    q[0] = quant_int(dc_coef_from_sum((int) d[0]), chroma ? 17 : 16);
    for (int i = 0; i < 64; i++) chunk[i] = q[scan_tab(i)];
    *amb_out = amb;
}

int emul_rle(const int* chunk, int prev_dc, uint8_t* out) {
    uint64_t nz = 0;
    for (int i = 1; i < 64; i++) if (chunk[i]) nz |= 1ull << i;
    uint32_t words[32];
    auto coef = [&](int i) { return chunk[i]; };
    auto put = [&](int w, uint32_t be) { words[w] = __builtin_bswap32(be); };
    int n = rle_chunk(chunk[0], prev_dc, nz, coef, put);
    memcpy(out, words, n);
    return n;
}

void emul_rgb2ycc_n(const uint8_t* rgb, long n, uint8_t* ycc) {
    for (long i = 0; i < n; i++) {
        int r = rgb[3 * i], g = rgb[3 * i + 1], b = rgb[3 * i + 2];
        ycc[3 * i] = (uint8_t) luma_of(r, g, b);
        ycc[3 * i + 1] = (uint8_t) cb_of(r, g, b);
        ycc[3 * i + 2] = (uint8_t) cr_of(r, g, b);
    }
}

void emul_ycc2rgb_n(const uint8_t* ycc, long n, uint8_t* rgb) {
    for (long i = 0; i < n; i++) {
        int y = ycc[3 * i], cb = ycc[3 * i + 1], cr = ycc[3 * i + 2];
        rgb[3 * i] = (uint8_t) red_of(y, cr);
        rgb[3 * i + 1] = (uint8_t) green_of(y, cb, cr);
        rgb[3 * i + 2] = (uint8_t) blue_of(y, cb);
    }
}

int emul_dc_from_sum(int S, int chroma) { return quant_int(dc_coef_from_sum(S), chroma ? 17 : 16); }
int emul_quant_int(int c, int Q) { return quant_int(c, Q); }
void emul_cos_bits(uint32_t* out) { for (int i = 0; i < 64; i++) out[i] = cos_bits(i); }

}
