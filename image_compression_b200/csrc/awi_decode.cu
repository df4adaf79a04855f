// awi_decode.cu — AWI payload -> pixels for sm_100a: the inverse of awi_encode.cu.
//
// The reference decoder (awi.c:184-217) is serial twice over: a chunk's start is only
// known once the previous chunk has been parsed (rle.c:134-167: no index, no markers,
// EOB is not byte aligned inside a chunk) and each DC is a running sum (rle.c:123-128).
// Here both become scans:
//
//   1  dec_scan_segments   per 2 KB segment: the byte offsets that can follow an end-of-block (8 zero
//                          bits ending in the previous byte) are collected, each speculatively
//                          parses the chunk that would start there -> L[p]; pointer jumping over
//                          that list gives, for each of the 184 possible entry offsets,
//                          entry -> (exit, #chunks)
//   2  dec_compose         32-ary tree of those maps (function composition is associative)
//   3  dec_root_expand,    top-down: true entry offset + first chunk index of every node (the root check and
//      dec_expand          the small upper levels in one CTA, the segment level in parallel)
//   4  dec_emit_offsets    per segment (one lane each): walk from the true entry, write chunk_off[]
//   5  dec_dc_*            per-stream (chunk index mod 6) inclusive scan of the DC diffs
//   6  dec_reconstruct     8 macroblocks per CTA: parse -> dequantise -> IDCT in the
//                          reference's exact fp64 operation order (dct.c:64-78, zero terms
//                          skipped) -> 4:2:0 replicate -> inverse colour -> BMP rows / RGB24
//
// Steps 1-4 are skipped when the encoder's chunk_off side-info is supplied.
#include <cuda_runtime.h>
#include <stdint.h>
#include "codec_math.cuh"
#include "awi_kernels.h"
#include <mutex>

namespace awi {

constexpr int SEG = 2048;            // payload bytes per discovery segment
constexpr int MAXCH = 184;           // longest legal chunk: 12 + 63*(8+15) + 8 bits
constexpr int SEG_STAGE = SEG + 208; // segment + longest chunk + read-ahead
constexpr int FAN = 32;              // children per node of the composition tree
constexpr int DCT_MB = 256;          // macroblocks per DC-scan tile
#ifndef REC_MB_N
#define REC_MB_N 8
#endif
constexpr int REC_MB = REC_MB_N;     // macroblocks per reconstruct CTA
#ifndef REC_ROWS
#define REC_ROWS 2                       // image rows per IDCT thread (8/REC_ROWS threads per block)
#endif
constexpr int REC_TPB = 8 / REC_ROWS;     // IDCT threads per 8x8 block
constexpr int REC_THREADS = REC_MB * 6 * REC_TPB;

// map entry: bit 63 valid | bits 8..47 chunk count | bits 0..7 exit offset into the next node.
// Without bit 63 the chain died on the way (no legal chunk where one had to start, or the payload ended inside
// one); the count then says how many whole chunks it got through, which is what decides whether a stream with
// bytes after its last chunk is still complete (the reference stops reading after the last chunk, awi.c:184-217).
constexpr unsigned long long ENT_OK = 1ull << 63;
__device__ __forceinline__ unsigned long long ent_make(unsigned long long cnt, int exit) {
    return ENT_OK | (cnt << 8) | (unsigned long long) exit;
}
__device__ __forceinline__ unsigned long long ent_dead(unsigned long long cnt) { return cnt << 8; }
__device__ __forceinline__ unsigned long long ent_cnt(unsigned long long e) { return (e >> 8) & ((1ull << 40) - 1); }
__device__ __forceinline__ int ent_exit(unsigned long long e) { return (int) (e & 0xff); }

// ---------------------------------------------------------------- 1: segments
// Chunk length when parsing from byte s[0] (rle.c:113-167 without producing values);
// 0 = not a legal chunk (more than 64 coefficients, or longer than MAXCH bytes).
// `words` is the 16-byte aligned stage, `p` the byte position of the chunk.  A 32-bit window of the stream serves two
// symbols (a symbol and its value are at most 23 bits, the next symbol's 8 bits still lie inside), so the loop runs
// once per two symbols and reads words instead of bytes (noise content: 8192x8192 decode 1.63 -> 1.59 ms).
__device__ __forceinline__ int speculative_len_words(const uint32_t* words, int p) {
    uint32_t ab = (uint32_t) p * 8u + 12u;                      // absolute bit position in the stage
    const uint32_t start = (uint32_t) p * 8u;
    int cnt = 1;
    for (int it = 0; it < 32; it++) {
        const uint32_t w0 = __byte_perm(words[ab >> 5], 0, 0x0123), w1 = __byte_perm(words[(ab >> 5) + 1], 0, 0x0123);
        const uint32_t win = __funnelshift_l(w1, w0, ab & 31u);   // 32 stream bits from `ab`
        const uint32_t sym1 = win >> 24;
        if (sym1 == 0) return (int) ((ab - start + 8u + 7u) >> 3);
        const uint32_t adv1 = 8u + (sym1 & 15u);
        const uint32_t sym2 = (win << adv1) >> 24;
        cnt += (int) (sym1 >> 4) + 1;
        if (sym2 == 0) return cnt > 64 ? 0 : (int) ((ab - start + adv1 + 8u + 7u) >> 3);
        cnt += (int) (sym2 >> 4) + 1;
        ab += adv1 + 8u + (sym2 & 15u);
        if (cnt > 64) return 0;
    }
    return 0;
}

__global__ void __launch_bounds__(256) dec_scan_segments(const uint8_t* __restrict__ payload, unsigned long long P,
                                                         uint8_t* __restrict__ Lg, unsigned long long* __restrict__ table) {
    pdl_wait();
    pdl_trigger();
    __shared__ __align__(16) uint8_t s[SEG_STAGE];
    __shared__ uint16_t idx[SEG];            // position -> index in the candidate list (NONE: not a candidate)
    __shared__ uint16_t cand_pos[SEG];       // candidate list, ascending positions
    __shared__ uint8_t cand_len[SEG];        // chunk length parsed from each candidate (0: no legal chunk there)
    __shared__ unsigned jump[SEG];           // per candidate: (where its chain has got to, chunks counted)
    __shared__ int warp_cnt[8];
    constexpr unsigned NONE = 0xffffu, DEAD = 0xffffu, EXITED = 0x8000u;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned long long seg_start = (unsigned long long) blockIdx.x * SEG;
    if ((reinterpret_cast<uintptr_t>(payload) & 15) == 0 && seg_start + SEG_STAGE <= P) {      // whole stage inside the payload
        static_assert(SEG_STAGE % 16 == 0, "stage is copied in 16-byte pieces");
        const uint4* src = reinterpret_cast<const uint4*>(payload + seg_start);
        for (int i = tid; i < SEG_STAGE / 16; i += 256) reinterpret_cast<uint4*>(s)[i] = __ldg(src + i);
    } else {
        for (int i = tid; i < SEG_STAGE; i += 256) {
            unsigned long long g = seg_start + i;
            s[i] = g < P ? payload[g] : 0;
        }
    }
    __syncthreads();
    // A chunk can only start where the one before it ended, and every chunk ends with the 8 zero bits of the
    // end-of-block symbol (rle.c:106) followed by at most 7 padding bits that the parser skips unread
    // (bitrw.c:72-76): so the two bytes in front of a chunk start hold 8 consecutive zero bits that end inside
    // the last one.  Positions without that can never be reached by a parse (about two in three on photo-like
    // streams, four in five on noise); the others are collected, in order, and everything after this works on
    // that list.  Thread t looks at positions 8t .. 8t+7.  (L[] is only ever read at positions a chain visits,
    // and those are candidates: the others are not written at all.)
    unsigned flags = 0;
    {
        const int p0 = 8 * tid;
        unsigned b2 = 0, b1 = 0;                                         // the two bytes in front of position p
        if (p0 >= 2) { b2 = s[p0 - 2]; b1 = s[p0 - 1]; }
        else if (seg_start >= 2) { b2 = payload[seg_start - 2]; b1 = payload[seg_start - 1]; }      // thread 0 of a later segment
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const unsigned long long g = seg_start + p0 + k;
            bool cand = false;
            if (g < P) {
                if (g == 0) {
                    cand = true;
                } else if (g >= 3) {                                     // a chunk is at least 3 bytes
                    const unsigned z = ~((b2 << 8) | b1) & 0xffffu;      // ones where the stream has zeros
                    unsigned r = z & (z >> 1);
                    r &= r >> 2;
                    r &= r >> 4;                                         // bit i: window bits i..i+7 are all zero
                    cand = (r & 0xffu) != 0;                             // the run ends inside the last byte
                }
            }
            flags |= (cand ? 1u : 0u) << k;
            b2 = b1;
            b1 = s[p0 + k];
        }
    }
    int n_cand;
    {
        const int mine = __popc(flags);
        int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) warp_cnt[warp] = incl;
        __syncthreads();
        int base = 0, total = 0;
#pragma unroll
        for (int w = 0; w < 8; w++) { const int t = warp_cnt[w]; if (w < warp) base += t; total += t; }
        n_cand = total;
        int at = base + incl - mine;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const bool cand = (flags >> k) & 1u;
            idx[8 * tid + k] = cand ? (uint16_t) at : (uint16_t) NONE;
            if (cand) cand_pos[at++] = (uint16_t) (8 * tid + k);
        }
    }
    __syncthreads();
    for (int i = tid; i < n_cand; i += 256) {
        const int p = cand_pos[i];
        const unsigned long long g = seg_start + p;
        int len = speculative_len_words(reinterpret_cast<const uint32_t*>(s), p);
        if (g + len > P) len = 0;
        cand_len[i] = (uint8_t) len;
        Lg[g] = (uint8_t) len;
    }
    __syncthreads();
    // entry offset -> (exit offset into the next segment, chunks passed) for each of the 184 possible entry
    // offsets, by pointer jumping over the candidates: jump[i] = (where the chain from candidate i has got to,
    // chunks counted), one 32-bit word each so a pair is always read consistently; 10 doublings cover the
    // longest chain (2048/3 chunks).  In-place updates may see a neighbour's old or new word: both are valid jumps
    // (a word is either (live index, chunks to get there) or final).
    const long long remain = (long long) (P - seg_start);           // bytes of payload from seg_start on
    const int lim = remain < SEG ? (int) remain : SEG;              // in the last segment a legal chain ends exactly at lim
    for (int i = tid; i < n_cand; i += 256) {
        unsigned w = DEAD;                                           // (count 0: no chunk starts here)
        const int l = cand_len[i], p = cand_pos[i];
        if (l != 0 && p < lim) {
            const int q = p + l;                                     // q <= lim in the last segment (longer chunks were zeroed)
            const unsigned to = q < lim ? (unsigned) idx[q] : EXITED + (unsigned) (q - lim);      // NONE == DEAD: the chain stops at q
            w = to | (1u << 16);                                     // one chunk passed, even if nothing can follow it
        }
        jump[i] = w;
    }
    __syncthreads();
    // Four doublings (chains advance 16 chunks per jump), not the ten that the longest chain would need: the 184 entry
    // threads walk what is left (a dozen dependent reads).  Measured at 8192x8192: 10 rounds 0.463 ms, 6: 0.457,
    // 4: 0.455, 3: 0.457.
    constexpr int SCAN_ROUNDS = 4;
    for (int round = 0; round < SCAN_ROUNDS; round++) {
        int moved = 0;
        for (int i = tid; i < n_cand; i += 256) {
            const unsigned w = jump[i], n = w & 0xffffu;
            if (n < EXITED) {
                const unsigned w2 = jump[n], n2 = w2 & 0xffffu;
                jump[i] = n2 | ((w >> 16) + (w2 >> 16)) << 16;       // a dead end keeps the number of chunks passed
                moved = 1;
            }
        }
        if (!__syncthreads_or(moved)) break;                         // every chain has left the segment (or died)
    }
    if (tid < MAXCH) {
        const int e = tid;
        unsigned long long r = 0ull;
        if (e == lim) {
            r = ent_make(0, 0);                                      // entering exactly at the payload end: nothing left
        } else if (e < lim && idx[e] != NONE) {
            unsigned w = jump[idx[e]], n = w & 0xffffu;
            // Fewer doubling rounds than the longest chain needs leave some chains inside the segment: each entry walks
            // the rest of its own chain in the jumps that are there (nobody writes them any more).
            while (n < EXITED) {
                const unsigned w2 = jump[n];
                n = w2 & 0xffffu;
                w = n | ((w >> 16) + (w2 >> 16)) << 16;
            }
            if (n != DEAD && n >= EXITED) r = ent_make(w >> 16, remain <= SEG ? 0 : (int) (n - EXITED));
            else if (n == DEAD) r = ent_dead(w >> 16);
        }
        table[(unsigned long long) blockIdx.x * MAXCH + tid] = r;
    }
}

// ---------------------------------------------------------------- 2: compose
// out[i][e] = children c0..c1 of node i applied in turn to entry e.  On the way, pre[c][e] records the state
// in front of child c (entry offset into it, chunks counted before it) so that the top-down pass needs one
// look-up per child instead of a chain of them.
// The chain through the children is a pointer chase; the children's tables are fetched into shared memory
// first (independent, coalesced loads) so that the 32 dependent steps cost shared-memory latency, not DRAM's.
__global__ void __launch_bounds__(192) dec_compose(const unsigned long long* __restrict__ in, long long n_in,
                                                   unsigned long long* __restrict__ out,
                                                   unsigned long long* __restrict__ pre) {
    pdl_wait();
    pdl_trigger();
    extern __shared__ unsigned long long child_tab[];               // FAN * MAXCH entries
    const long long c0 = (long long) blockIdx.x * FAN, c1 = c0 + FAN < n_in ? c0 + FAN : n_in;
    const int n = (int) (c1 - c0);
    for (int i = threadIdx.x; i < n * MAXCH; i += 192) child_tab[i] = in[c0 * MAXCH + i];
    __syncthreads();
    if (threadIdx.x >= MAXCH) return;
    int e = threadIdx.x;
    unsigned long long cnt = 0;
    bool ok = true;
    for (int c = 0; c < n; c++) {
        pre[(c0 + c) * MAXCH + threadIdx.x] = ok ? ent_make(cnt, e) : 0ull;
        if (ok) {
            unsigned long long t = child_tab[c * MAXCH + e];
            cnt += ent_cnt(t);                                       // (a dying child still counts the chunks it got through)
            if (!(t & ENT_OK)) ok = false;
            else e = ent_exit(t);
        }
    }
    out[(long long) blockIdx.x * MAXCH + threadIdx.x] = ok ? ent_make(cnt, e) : ent_dead(cnt);
}

// ---------------------------------------------------------------- 3: expand
// state[i] = (entry offset, first chunk index) of node i; one thread per child, no chains.
struct NodeState { unsigned long long first_chunk; int entry; int ok; };

__global__ void dec_expand(const unsigned long long* __restrict__ pre, long long n_child,
                           const NodeState* __restrict__ parent, NodeState* __restrict__ child) {
    pdl_wait();
    pdl_trigger();
    long long c = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_child) return;
    NodeState st = parent[c / FAN];
    if (st.ok) {
        const unsigned long long p = pre[c * MAXCH + st.entry];
        if (p & ENT_OK) { st.first_chunk += ent_cnt(p); st.entry = ent_exit(p); }
        else st.ok = 0;
    }
    child[c] = st;
}

// The root check and the expansion of the small upper levels (down to level 1) in one CTA: a launch per
// level is all latency there.
struct TopLevels {
    int levels;
    long long n_node[16];
    const unsigned long long* pre[16];            // pre[lv]: states in front of the children (level lv-1) of level lv
    NodeState* state[16];
};
__global__ void __launch_bounds__(256) dec_root_expand(const unsigned long long* __restrict__ top_table, TopLevels T,
                                                       unsigned long long expect_chunks, unsigned int* err) {
    pdl_wait();
    pdl_trigger();
    if (threadIdx.x == 0) {
        const unsigned long long t = top_table[0];
        NodeState st; st.first_chunk = 0; st.entry = 0; st.ok = 1;
        // complete = the chain from byte 0 gets through at least the expected number of chunks, whatever follows them
        if (ent_cnt(t) < expect_chunks) { st.ok = 0; atomicOr(err, 2u); }
        T.state[T.levels - 1][0] = st;
    }
    __syncthreads();
    for (int lv = T.levels - 1; lv >= 2; lv--) {
        for (long long c = threadIdx.x; c < T.n_node[lv - 1]; c += 256) {
            NodeState st = T.state[lv][c / FAN];
            if (st.ok) {
                const unsigned long long p = T.pre[lv][c * MAXCH + st.entry];
                if (p & ENT_OK) { st.first_chunk += ent_cnt(p); st.entry = ent_exit(p); }
                else st.ok = 0;
            }
            T.state[lv - 1][c] = st;
        }
        __syncthreads();                                  // this CTA's global writes, read by this CTA
    }
}

// ---------------------------------------------------------------- 4: offsets
// Walk every segment's chain from its true entry offset and write chunk_off[].  The walk is serial by
// nature, so a whole warp is spent on 32 of them at once (one segment per lane): the CTA first stages the
// 32 segments' length bytes in shared memory, then warp 0 walks.
constexpr int EMIT_SEGS = 32;
__global__ void __launch_bounds__(256) dec_emit_offsets(const uint8_t* __restrict__ Lg, unsigned long long P,
                                                        const NodeState* __restrict__ seg_state, long long n_seg,
                                                        uint32_t* __restrict__ chunk_off, unsigned long long n_chunks) {
    pdl_wait();
    pdl_trigger();
    extern __shared__ __align__(16) uint8_t Ls[];                    // EMIT_SEGS * SEG bytes
    const long long seg0 = (long long) blockIdx.x * EMIT_SEGS;
    const unsigned long long byte0 = (unsigned long long) seg0 * SEG;
    const unsigned long long span = min((unsigned long long) EMIT_SEGS * SEG, P - byte0);   // Lg holds P (+16 spare) bytes
    {
        const uint4* src = reinterpret_cast<const uint4*>(Lg + byte0);                  // byte0 is a multiple of SEG, Lg 256-aligned
        const int n16 = (int) ((span + 15) >> 4);
        for (int i = threadIdx.x; i < n16; i += 256) reinterpret_cast<uint4*>(Ls)[i] = __ldg(src + i);
    }
    __syncthreads();
    if (threadIdx.x >= 32) return;
    const long long seg = seg0 + threadIdx.x;
    if (seg >= n_seg) return;
    const NodeState st = seg_state[seg];
    if (!st.ok) return;
    const unsigned long long seg_start = (unsigned long long) seg * SEG;
    const long long remain = (long long) (P - seg_start);
    const int lim = remain < SEG ? (int) remain : SEG;
    const uint8_t* L = Ls + threadIdx.x * SEG;
    int pos = st.entry;
    unsigned long long c = st.first_chunk;
    bool alive = true;
    while (pos < lim && c < n_chunks) {
        chunk_off[c] = (uint32_t) (seg_start + pos);
        const int l = L[pos];
        if (l == 0) { alive = false; break; }
        pos += l;
        c++;
    }
    // where the last chunk ends (the payload end, or the first byte the reference would not read); the segment the
    // last chunk runs into may write the same value again
    if (alive && c == n_chunks) chunk_off[n_chunks] = (uint32_t) (seg_start + pos);
}

// ---------------------------------------------------------------- 5: DC scan
// rle.c:118-128: 12-bit field, raw > 2047 means raw - 4095.
__device__ __forceinline__ int dc_diff_at(const uint8_t* __restrict__ payload, uint32_t off, unsigned long long P) {
    if ((unsigned long long) off + 2 > P) off = (uint32_t) (P - 2);      // malformed side-info: stay in bounds
    int raw = ((int) payload[off] << 4) | ((int) payload[off + 1] >> 4);
    return raw > 2047 ? raw - 4095 : raw;
}

template <int MODE>   // 0: tile sums   1: apply (inclusive scan + tile base)   2: apply, adding up the sums of the tiles in front itself
__global__ void __launch_bounds__(DCT_MB) dec_dc_tile(const uint8_t* __restrict__ payload,
                                                      const uint32_t* __restrict__ chunk_off, long long n_mb,
                                                      int* __restrict__ tile_sum, const int* __restrict__ tile_base,
                                                      int* __restrict__ dcabs, unsigned long long P,
                                                      const unsigned int* __restrict__ err) {
    pdl_wait();
    pdl_trigger();
    __shared__ int wsum[DCT_MB / 32][6];
    __shared__ int bsum[DCT_MB / 32][6];
    if (*reinterpret_cast<const volatile unsigned int*>(err) != 0) return;     // boundaries unknown: nothing to scan
    const long long mb = (long long) blockIdx.x * DCT_MB + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (MODE == 2) {      // few tiles: every CTA sums what lies in front of it (tile_sum[] is `tile_base` here) instead of a scan launch
        int s[6] = { 0, 0, 0, 0, 0, 0 };
        for (long long t = threadIdx.x; t < (long long) blockIdx.x; t += DCT_MB)
#pragma unroll
            for (int k = 0; k < 6; k++) s[k] += tile_base[t * 6 + k];
#pragma unroll
        for (int k = 0; k < 6; k++) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s[k] += __shfl_xor_sync(0xffffffffu, s[k], o);
            if (lane == 0) bsum[warp][k] = s[k];
        }
    }
    int v[6];
#pragma unroll
    for (int k = 0; k < 6; k++) v[k] = mb < n_mb ? dc_diff_at(payload, chunk_off[mb * 6 + k], P) : 0;
#pragma unroll
    for (int k = 0; k < 6; k++) {
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, v[k], o);
            if (lane >= o) v[k] += t;
        }
        if (lane == 31) wsum[warp][k] = v[k];
    }
    __syncthreads();
    if (MODE == 0) {
        if (threadIdx.x < 6) {
            int s = 0;
            for (int w = 0; w < DCT_MB / 32; w++) s += wsum[w][threadIdx.x];
            tile_sum[(long long) blockIdx.x * 6 + threadIdx.x] = s;
        }
    } else {
        if (mb < n_mb) {
#pragma unroll
            for (int k = 0; k < 6; k++) {
                int base = 0;
                if (MODE == 2) { for (int w = 0; w < DCT_MB / 32; w++) base += bsum[w][k]; }
                else base = tile_base[(long long) blockIdx.x * 6 + k];
                for (int w = 0; w < warp; w++) base += wsum[w][k];
                dcabs[mb * 6 + k] = base + v[k];
            }
        }
    }
}

// exclusive scan of the tile sums (one CTA; each thread owns a contiguous run of tiles, the 1024 partial
// sums are scanned with shuffles)
__global__ void __launch_bounds__(1024) dec_dc_scan_tiles(const int* __restrict__ tile_sum, long long n_tiles,
                                                          int* __restrict__ tile_base) {
    pdl_wait();
    pdl_trigger();
    __shared__ int wtot[32][6];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long per = (n_tiles + 1023) / 1024;
    const long long t0 = min((long long) threadIdx.x * per, n_tiles), t1 = t0 + per < n_tiles ? t0 + per : n_tiles;
    int s[6] = { 0, 0, 0, 0, 0, 0 };
    for (long long t = t0; t < t1; t++)
#pragma unroll
        for (int k = 0; k < 6; k++) s[k] += tile_sum[t * 6 + k];
    int incl[6];
#pragma unroll
    for (int k = 0; k < 6; k++) {
        incl[k] = s[k];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl[k], o);
            if (lane >= o) incl[k] += t;
        }
        if (lane == 31) wtot[warp][k] = incl[k];
    }
    __syncthreads();
    if (warp == 0) {
#pragma unroll
        for (int k = 0; k < 6; k++) {
            const int w = wtot[lane][k];
            int wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, wi, o);
                if (lane >= o) wi += t;
            }
            wtot[lane][k] = wi - w;                       // exclusive over warps
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 6; k++) s[k] = wtot[warp][k] + incl[k] - s[k];          // exclusive prefix of my run
    for (long long t = t0; t < t1; t++)
#pragma unroll
        for (int k = 0; k < 6; k++) { tile_base[t * 6 + k] = s[k]; s[k] += tile_sum[t * 6 + k]; }
}

// ---------------------------------------------------------------- 6: reconstruct
struct __align__(16) RecSmem {
    int coef[REC_MB * 6][64];                 // dequantised coefficients, natural [u][v] order (only masked entries are valid)
    unsigned long long mask[REC_MB * 6];      // non-zero positions, natural order
    double cost[64];                          // cost[v*8 + x] = (double) cos_table[x][v]  (dct.c:14-23), frequency-major
    double cucv[64];                          // cu*cv per natural position (dct.c:71-72,75)
    uint32_t words[(REC_MB * 6 * MAXCH + 64) / 4];   // the tile's slice of the payload as big-endian words
    uint8_t yp[16][REC_MB * 16];
    uint8_t cbp[8][REC_MB * 8];
    uint8_t crp[8][REC_MB * 8];
    uint8_t scan[64];
    uint8_t q[2][64];
};

__device__ __noinline__ int g_exact_cold(unsigned int y, int cb, int cr) { return g_exact((int) y, cb, cr); }

__device__ __forceinline__ void store_px(const DecParams& P, int x, int y, int r, int g, int b) {
    if (P.rgb_topdown) {
        uint8_t* p = P.dst + ((long long) y * P.width + x) * 3;
        p[0] = (uint8_t) r; p[1] = (uint8_t) g; p[2] = (uint8_t) b;
    } else {
        uint8_t* p = P.dst + (long long) (P.height - 1 - y) * P.stride + 3ll * x;
        p[0] = (uint8_t) b; p[1] = (uint8_t) g; p[2] = (uint8_t) r;
    }
}

__global__ void __launch_bounds__(REC_THREADS) dec_reconstruct(DecParams P, const uint32_t* __restrict__ chunk_off,
                                                               const int* __restrict__ dcabs, unsigned int* err,
                                                               FastDiv div_tiles_per_row) {
    pdl_wait();
    pdl_trigger();
    extern __shared__ __align__(16) unsigned char smem_raw[];
    RecSmem& S = *reinterpret_cast<RecSmem*>(smem_raw);
    const int tid = threadIdx.x;
    const int tiles_per_row = (P.mbw + REC_MB - 1) / REC_MB;
    const int row_in = fast_div((int) blockIdx.x, div_tiles_per_row);
    const int my = P.mb_row0 + row_in;
    const int mx0 = ((int) blockIdx.x - row_in * tiles_per_row) * REC_MB;
    const int n_mb = min(REC_MB, P.mbw - mx0);
    const long long chunk0 = ((long long) my * P.mbw + mx0) * 6;
    const int n_ch = n_mb * 6;
    // (the flag and the offsets below are all requested before the first of them is looked at; the offsets array
    // exists whatever the flag says)
    const unsigned int rejected = *reinterpret_cast<volatile unsigned int*>(err);

    // the tile's chunks are contiguous in the stream: [b0, b1)
    const uint32_t b0 = chunk_off[chunk0], b1_raw = chunk_off[chunk0 + n_ch];
    uint32_t my_off = 0;
    int my_dc = 0;
    if (tid < n_ch) { my_off = chunk_off[chunk0 + tid]; my_dc = dcabs[chunk0 + tid]; }      // in flight during staging
    if (rejected != 0) return;                                                // an earlier stage rejected the stream
    uint32_t b1 = b1_raw;
    // With the encoder's side-info the end of the last chunk is not known, only the end of the payload: whatever
    // follows the last chunk is ignored (as the reference does, awi.c:184-217), so the last tile's slice is cut
    // to the longest it can legally be.  Anywhere else an over-long slice means inconsistent offsets.
    const bool last_tile = chunk0 + n_ch == (long long) P.mbw * P.mbh * 6;
    if (last_tile && b1 >= b0 && b1 - b0 > (uint32_t) (REC_MB * 6 * MAXCH)) b1 = b0 + (uint32_t) (REC_MB * 6 * MAXCH);
    if (b1 < b0 || b1 > P.payload_len || b1 - b0 > (uint32_t) (REC_MB * 6 * MAXCH)) { if (tid == 0) atomicOr(err, 4u); return; }
    if (tid < 64) {
        const double c = (double) __uint_as_float(cos_bits_rt(tid));          // cos_table[x = tid>>3][v = tid&7]
        S.cost[(tid & 7) * 8 + (tid >> 3)] = c;
        const double inv_sqrt2 = 0.70710678118654746;
        S.cucv[tid] = __dmul_rn((tid >> 3) == 0 ? inv_sqrt2 : 1.0, (tid & 7) == 0 ? inv_sqrt2 : 1.0);
        S.scan[tid] = (uint8_t) scan_rt(tid);
        S.q[0][tid] = (uint8_t) qluma_rt(tid);
        S.q[1][tid] = (uint8_t) qchroma_rt(tid);
    }
    // stage [a0, b1 + 8) as big-endian words, a0 = b0 rounded down to 4
    const uint32_t a0 = b0 & ~3u;
    const uint32_t nbytes = min(b1 - b0, (uint32_t) (REC_MB * 6 * MAXCH));
    {
        const uint32_t n_words = (b0 - a0 + nbytes + 8 + 3) >> 2;
        const uint32_t* src = reinterpret_cast<const uint32_t*>(P.payload + a0);
        const bool aligned = (reinterpret_cast<uintptr_t>(P.payload) & 3) == 0;
        const uint32_t safe_words = (uint32_t) ((P.payload_len - a0) >> 2);   // whole words inside the buffer
        for (uint32_t i = tid; i < n_words; i += REC_THREADS) {
            uint32_t w = 0;
            if (aligned && i < safe_words) {
                w = __ldg(src + i);
            } else {
                for (int k = 0; k < 4; k++) {
                    unsigned long long g = (unsigned long long) a0 + 4ull * i + k;
                    if (g < P.payload_len) w |= (uint32_t) P.payload[g] << (8 * k);
                }
            }
            S.words[i] = __byte_perm(w, 0, 0x0123);
        }
    }
    __syncthreads();

    // ---- parse: one thread per chunk (rle.c:113-170), dequantise (dct.c:128-145) and
    //      scatter to natural order (dct.c:171-181)
    if (tid < n_ch) {
        const int type = tid % 6;
        const uint8_t* qt = S.q[type >= 4 ? 1 : 0];
        const uint32_t base_bit = (my_off - a0) * 8;                          // chunk start, in bits from S.words[0]
        const int limit_bits = (int) (b1 - my_off) * 8;
        int* out = S.coef[tid];
        unsigned long long mask = 0;
        const int dc = my_dc * qt[0];
        out[0] = dc;
        if (dc != 0) mask |= 1ull;
        int bit = 12, idx = 1;
        bool bad = my_off < b0 || my_off > b1;
        while (!bad) {
            if (bit + 8 > limit_bits) { bad = true; break; }
            const uint32_t ab = base_bit + (uint32_t) bit;
            const uint32_t win = __funnelshift_l(S.words[(ab >> 5) + 1], S.words[ab >> 5], ab & 31);   // 32 stream bits from `bit`
            const int sym = win >> 24;
            if (sym == 0) break;
            const int run = sym >> 4, size = sym & 15;
            idx += run;
            if (idx > 63) { bad = true; break; }
            if (size > 0) {
                const int val = (int) ((win << 8) >> (32 - size));
                const int v = (val & (1 << (size - 1))) ? val : val - (1 << size) + 1;      // rle.c:152-155
                const int k = S.scan[idx];
                const int c = v * qt[k];
                out[k] = c;
                if (c != 0) mask |= 1ull << k;
            }
            idx++;
            bit += 8 + size;
        }
        if (bad) { atomicOr(err, 4u); mask = 0; }
        S.mask[tid] = mask;
    }
    __syncthreads();

    // ---- IDCT: REC_TPB threads per block, REC_ROWS output rows each.  dct.c:64-78: for u, for v:
    //      sum += (((cu*cv)*coef)*cos[y][u])*cos[x][v], all fp64, strictly in that order.
    {
        const int blk = tid / REC_TPB, y0 = (tid % REC_TPB) * REC_ROWS;
        if (blk < n_ch) {
            double sum[REC_ROWS][8];
            unsigned long long m = S.mask[blk];
            const int* cf = S.coef[blk];
            // The DC term is the same for all 64 samples: cos_table[.][0] is 1.0f, so ((cu*cv*coef) * 1.0) * 1.0 is the
            // first product itself and 0.0 + it is exact: the sums start from it instead of taking a turn of the loop.
            double first = 0.0;
            if (m & 1ull) { first = __dmul_rn(S.cucv[0], (double) cf[0]); m &= ~1ull; }
#pragma unroll
            for (int r = 0; r < REC_ROWS; r++)
#pragma unroll
                for (int x = 0; x < 8; x++) sum[r][x] = first;
            // the mask as two words, the lower non-empty one served first: a dozen instructions of loop control per
            // coefficient instead of the two dozen that 64-bit find-first-set / clear-lowest / test compile to
            uint32_t lo = (uint32_t) m, hi = (uint32_t) (m >> 32);
            while ((lo | hi) != 0u) {
                const bool low = lo != 0u;
                const uint32_t t = low ? lo : hi, rest = t & (t - 1u);
                const int k = (__ffs((int) t) - 1) + (low ? 0 : 32);
                lo = low ? rest : 0u;
                hi = low ? hi : rest;
                const int u = k >> 3, v = k & 7;
                const double a = __dmul_rn(S.cucv[k], (double) cf[k]);
                double b[REC_ROWS];
#pragma unroll
                for (int r = 0; r < REC_ROWS; r++) b[r] = __dmul_rn(a, S.cost[u * 8 + y0 + r]);        // cos_table[y][u]
                const double2* cx = reinterpret_cast<const double2*>(&S.cost[v * 8]);          // cos_table[x][v], x = 0..7
#pragma unroll
                for (int x2 = 0; x2 < 4; x2++) {
                    const double2 c = cx[x2];
#pragma unroll
                    for (int r = 0; r < REC_ROWS; r++) {
                        sum[r][2 * x2] = __dadd_rn(sum[r][2 * x2], __dmul_rn(b[r], c.x));
                        sum[r][2 * x2 + 1] = __dadd_rn(sum[r][2 * x2 + 1], __dmul_rn(b[r], c.y));
                    }
                }
            }
            const int mb = blk / 6, type = blk - 6 * mb;
#pragma unroll
            for (int r = 0; r < REC_ROWS; r++) {
                const int y = y0 + r;
                uint32_t px[8];
#pragma unroll
                for (int x = 0; x < 8; x++) px[x] = (uint32_t) clamp_trunc(__dadd_rn(__dmul_rn(0.25, sum[r][x]), 128.0));
                uint8_t* dstp;
                if (type < 4) dstp = &S.yp[((type >> 1) << 3) + y][mb * 16 + ((type & 1) << 3)];
                else if (type == 4) dstp = &S.cbp[y][mb * 8];
                else dstp = &S.crp[y][mb * 8];
                uint2 w;
                w.x = px[0] | (px[1] << 8) | (px[2] << 16) | (px[3] << 24);
                w.y = px[4] | (px[5] << 8) | (px[6] << 16) | (px[7] << 24);
                *reinterpret_cast<uint2*>(dstp) = w;
            }
        }
    }
    __syncthreads();

    // ---- 4:2:0 replicate (color.c:88-105) + inverse colour (color.c:45-67) + store with
    //      cropping (bmp.c:250, :304-308).  4 consecutive pixels per thread.
    {
        const int quads_per_row = REC_MB * 4;
        for (int item = tid; item < 16 * quads_per_row; item += REC_THREADS) {
            int r = item / quads_per_row, xq = item - r * quads_per_row;
            int xl = xq * 4;
            int y = my * 16 + r, x = mx0 * 16 + xl;
            if (y >= P.height || xl >= n_mb * 16) continue;
            uint32_t yw = *reinterpret_cast<const uint32_t*>(&S.yp[r][xl]);
            uint16_t cbw = *reinterpret_cast<const uint16_t*>(&S.cbp[r >> 1][xl >> 1]);
            uint16_t crw = *reinterpret_cast<const uint16_t*>(&S.crp[r >> 1][xl >> 1]);
            int R[4], G[4], B[4];
#pragma unroll
            for (int c2 = 0; c2 < 2; c2++) {                      // two chroma samples, two pixels each
                const int cb = (cbw >> (8 * c2)) & 0xff, cr = (crw >> (8 * c2)) & 0xff;
                const ChromaTerms t = chroma_terms(cb, cr);
#pragma unroll
                for (int i = 2 * c2; i < 2 * c2 + 2; i++) {
                    const int yy = (yw >> (8 * i)) & 0xff;
                    R[i] = red_from(yy, t); B[i] = blue_from(yy, t);
                    G[i] = clamp255(yy + t.g_add);
                }
                // The two (Cb, Cr) pairs whose green term is a non-zero integer take the literal fp64 expression
                // (codec_math.cuh).  Kept out of line: inlined, the compiler turns the branch into predicated code and every
                // pixel pays for the conversions and multiplications of a path that almost none takes (8192x8192: 0.454 -> 0.441 ms).
                if (t.g_on_integer) {
#pragma unroll
                    for (int i = 2 * c2; i < 2 * c2 + 2; i++) G[i] = g_exact_cold((yw >> (8 * i)) & 0xff, cb, cr);
                }
            }
            uint8_t* p;
            uint32_t w0, w1, w2;
            if (P.rgb_topdown) {
                p = P.dst + ((long long) y * P.width + x) * 3;
                w0 = R[0] | (G[0] << 8) | (B[0] << 16) | ((uint32_t) R[1] << 24);
                w1 = G[1] | (B[1] << 8) | (R[2] << 16) | ((uint32_t) G[2] << 24);
                w2 = B[2] | (R[3] << 8) | (G[3] << 16) | ((uint32_t) B[3] << 24);
            } else {
                p = P.dst + (long long) (P.height - 1 - y) * P.stride + 3ll * x;
                w0 = B[0] | (G[0] << 8) | (R[0] << 16) | ((uint32_t) B[1] << 24);
                w1 = G[1] | (R[1] << 8) | (B[2] << 16) | ((uint32_t) G[2] << 24);
                w2 = R[2] | (B[3] << 8) | (G[3] << 16) | ((uint32_t) R[3] << 24);
            }
            if (x + 4 <= P.width && (reinterpret_cast<uintptr_t>(p) & 3) == 0) {
                uint32_t* pw = reinterpret_cast<uint32_t*>(p);
                pw[0] = w0; pw[1] = w1; pw[2] = w2;
            } else {
#pragma unroll
                for (int i = 0; i < 4; i++)
                    if (x + i < P.width) store_px(P, x + i, y, R[i], G[i], B[i]);
            }
        }
    }
}

__global__ void dec_set_word(uint32_t* p, uint32_t v) {
    pdl_wait();
    pdl_trigger(); *p = v; }

// ---------------------------------------------------------------- host side
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct DecLayout {
    long long n_seg;
    int levels;                       // tables[0] = segments ... tables[levels-1] has one node
    long long n_node[16];
    size_t off_table[16], off_state[16], off_pre[16];   // off_pre[lv]: states in front of the children (level lv-1 nodes) of level lv
    size_t off_L, off_chunk_off, off_dcabs, off_tile_sum, off_tile_base, total;
    long long n_dct;
};

static DecLayout dec_layout(unsigned long long P, long long n_chunks) {
    DecLayout Lo{};
    Lo.n_seg = (long long) ((P + SEG - 1) / SEG);
    if (Lo.n_seg < 1) Lo.n_seg = 1;
    size_t off = 0;
    long long n = Lo.n_seg;
    int lv = 0;
    while (true) {
        Lo.n_node[lv] = n;
        Lo.off_table[lv] = off; off = align_up(off + (size_t) n * MAXCH * 8, 256);
        Lo.off_state[lv] = off; off = align_up(off + (size_t) n * sizeof(NodeState), 256);
        Lo.off_pre[lv + 1] = off; off = align_up(off + (size_t) n * MAXCH * 8, 256);
        lv++;
        if (n == 1) break;
        n = (n + FAN - 1) / FAN;
    }
    Lo.levels = lv;
    Lo.off_L = off; off = align_up(off + (size_t) P + 16, 256);
    Lo.off_chunk_off = off; off = align_up(off + (size_t) (n_chunks + 1) * 4, 256);
    Lo.off_dcabs = off; off = align_up(off + (size_t) n_chunks * 4, 256);
    Lo.n_dct = (n_chunks / 6 + DCT_MB - 1) / DCT_MB;
    Lo.off_tile_sum = off; off = align_up(off + (size_t) Lo.n_dct * 6 * 4, 256);
    Lo.off_tile_base = off; off = align_up(off + (size_t) Lo.n_dct * 6 * 4, 256);
    Lo.total = off;
    return Lo;
}

size_t dec_workspace_bytes(unsigned long long payload_len, long long n_chunks) {
    return dec_layout(payload_len, n_chunks).total;
}

cudaError_t launch_decode(const DecParams& P, void* workspace, size_t workspace_bytes,
                          const uint32_t* chunk_off_in, bool reuse_index, unsigned int* err, cudaStream_t stream) {
    const long long n_mb = (long long) P.mbw * P.mbh;
    const long long n_chunks = n_mb * 6;
    DecLayout Lo = dec_layout(P.payload_len, n_chunks);
    if (workspace_bytes < Lo.total) return cudaErrorInvalidValue;
    {   // kernels with more than 48 KB of dynamic shared memory have to opt in, once per device
        static bool attr_set[64] = { false };
        static std::mutex attr_mu;                   // contexts of several devices may launch from different host threads
        std::lock_guard<std::mutex> attr_lock(attr_mu);
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev < 0 || dev >= 64 || !attr_set[dev]) {
            cudaError_t e = cudaFuncSetAttribute(dec_reconstruct, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sizeof(RecSmem));
            if (e == cudaSuccess) e = cudaFuncSetAttribute(dec_emit_offsets, cudaFuncAttributeMaxDynamicSharedMemorySize, EMIT_SEGS * SEG);
            if (e != cudaSuccess) return e;
            if (dev >= 0 && dev < 64) attr_set[dev] = true;
        }
    }
    uint8_t* ws = static_cast<uint8_t*>(workspace);
    uint32_t* chunk_off = reinterpret_cast<uint32_t*>(ws + Lo.off_chunk_off);

    int* dcabs = reinterpret_cast<int*>(ws + Lo.off_dcabs);
    int* tile_sum = reinterpret_cast<int*>(ws + Lo.off_tile_sum);
    int* tile_base = reinterpret_cast<int*>(ws + Lo.off_tile_base);
    if (reuse_index) {
        // chunk offsets and DC values of this payload are already in the workspace (an earlier call on
        // the same stream): only reconstruct another range of macroblock rows
    } else if (chunk_off_in == nullptr) {
        uint8_t* Lg = ws + Lo.off_L;
        auto table = [&](int lv) { return reinterpret_cast<unsigned long long*>(ws + Lo.off_table[lv]); };
        auto state = [&](int lv) { return reinterpret_cast<NodeState*>(ws + Lo.off_state[lv]); };
        launch_pdl(dec_scan_segments, (unsigned) Lo.n_seg, 256, 0, stream, P.payload, P.payload_len, Lg, table(0));
        auto pre = [&](int lv) { return reinterpret_cast<unsigned long long*>(ws + Lo.off_pre[lv]); };
        for (int lv = 1; lv < Lo.levels; lv++)
            launch_pdl(dec_compose, (unsigned) Lo.n_node[lv], 192, FAN * MAXCH * 8, stream, table(lv - 1), Lo.n_node[lv - 1], table(lv), pre(lv));
        TopLevels T{};
        T.levels = Lo.levels;
        for (int lv = 0; lv < Lo.levels; lv++) { T.n_node[lv] = Lo.n_node[lv]; T.pre[lv] = pre(lv); T.state[lv] = state(lv); }
        launch_pdl(dec_root_expand, 1, 256, 0, stream, table(Lo.levels - 1), T, (unsigned long long) n_chunks, err);
        if (Lo.levels >= 2)
            launch_pdl(dec_expand, (unsigned) ((Lo.n_node[0] + 127) / 128), 128, 0, stream, pre(1), Lo.n_node[0], state(1), state(0));
        launch_pdl(dec_emit_offsets, (unsigned) ((Lo.n_seg + EMIT_SEGS - 1) / EMIT_SEGS), 256, EMIT_SEGS * SEG, stream,
            Lg, P.payload_len, state(0), Lo.n_seg, chunk_off, (unsigned long long) n_chunks);
    } else {
        cudaError_t e = cudaMemcpyAsync(chunk_off, chunk_off_in, (size_t) n_chunks * 4, cudaMemcpyDeviceToDevice, stream);
        if (e != cudaSuccess) return e;
        launch_pdl(dec_set_word, 1, 1, 0, stream, chunk_off + n_chunks, (uint32_t) P.payload_len);     // the value rides in the launch, no host buffer to outlive
    }

    if (!reuse_index) {
    launch_pdl(dec_dc_tile<0>, (unsigned) Lo.n_dct, DCT_MB, 0, stream, P.payload, chunk_off, n_mb, tile_sum, nullptr, nullptr, P.payload_len, err);
    if (Lo.n_dct <= 1024) {     // up to 8192x8192: the scan of the tile sums is folded into the second pass
        launch_pdl(dec_dc_tile<2>, (unsigned) Lo.n_dct, DCT_MB, 0, stream, P.payload, chunk_off, n_mb, nullptr, tile_sum, dcabs, P.payload_len, err);
    } else {
        launch_pdl(dec_dc_scan_tiles, 1, 1024, 0, stream, tile_sum, Lo.n_dct, tile_base);
        launch_pdl(dec_dc_tile<1>, (unsigned) Lo.n_dct, DCT_MB, 0, stream, P.payload, chunk_off, n_mb, nullptr, tile_base, dcabs, P.payload_len, err);
    }
    }

    const int tiles_per_row = (P.mbw + REC_MB - 1) / REC_MB;
    const int rows = P.mb_row1 - P.mb_row0;
    if (rows > 0)
        launch_pdl(dec_reconstruct, (unsigned) (tiles_per_row * rows), REC_THREADS, sizeof(RecSmem), stream, P, chunk_off, dcabs, err,
                   make_fastdiv((uint32_t) tiles_per_row));
    return cudaGetLastError();
}

}  // namespace awi
