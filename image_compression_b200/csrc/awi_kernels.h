// awi_kernels.h — internal interface between the C-ABI layer (gpu_codec.cu) and the
// kernel translation units.  Not installed; the public surface is include/gpu_codec.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "codec_tables.h"
#include "codec_math.cuh"

namespace awi {

#ifndef ENC_TILE_MB_N
#define ENC_TILE_MB_N 32
#endif
constexpr int ENC_TILE_MB = ENC_TILE_MB_N;      // macroblocks per encode tile (one CTA)

// Run-time indexed copies of the generated tables (device global memory; the hot loops
// use the constexpr accessors in codec_math.cuh, which fold into immediates).
static __device__ const int d_qluma[64] = AWI_QLUMA;
static __device__ const int d_qchroma[64] = AWI_QCHROMA;
static __device__ const int d_unscan[64] = AWI_UNSCAN;
static __device__ const int d_scan[64] = AWI_SCAN;
static __device__ const uint32_t d_cos_bits[64] = AWI_COS_BITS;
// fast-path quantiser constants, [0] luma, [1] chroma (codec_math.cuh quant_flag)
static __device__ const float d_gk[2][64] = { AWI_GK_LUMA, AWI_GK_CHROMA };
static __device__ const float d_gc[2][64] = { AWI_GC_LUMA, AWI_GC_CHROMA };
static __device__ const float d_gt[2][64] = { AWI_GT_LUMA, AWI_GT_CHROMA };
__device__ __forceinline__ int qluma_rt(int k) { return d_qluma[k]; }
__device__ __forceinline__ int qchroma_rt(int k) { return d_qchroma[k]; }
__device__ __forceinline__ int unscan_rt(int k) { return d_unscan[k]; }
__device__ __forceinline__ int scan_rt(int i) { return d_scan[i]; }
__device__ __forceinline__ uint32_t cos_bits_rt(int i) { return d_cos_bits[i]; }

// ---- programmatic dependent launch.  Every kernel of this library starts with pdl_wait() (the grid in front of it in
// the stream has completed and its writes are visible) followed by pdl_trigger() (the grid behind it may be placed
// on the SMs now and park at its own pdl_wait()), and is launched through launch_pdl(): launch latency and the
// ramp-down of one grid overlap the ramp-up of the next, which is most of what a short encode or decode step costs
// beyond its kernels.  GPU_CODEC_NO_PDL=1 (measurement knob) launches them the ordinary way.
#if defined(__CUDACC__)
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
bool pdl_enabled();
template <class... KArgs, class... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), unsigned grid, unsigned block, size_t smem, cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

// ---- encode
struct EncParams {
    const uint8_t* rows;          // address of bottom-up row 0 of the FULL image (may lie outside the
                                  // caller's window; only rows the tile range needs are dereferenced)
    long long stride;             // bytes between rows
    int width, height;            // full image
    int align;                    // 16, 4 or 1: guaranteed alignment of rows and stride
    int mbw;                      // macroblocks per row
    int mb0, mb1;                 // raster macroblock range [mb0, mb1) to encode (of every frame of a batch)
    int tiles_per_frame;          // batches: tiles per frame (0 = single frame); the grid covers frames x tiles
    FastDiv div_mbw, div_tpf;     // filled by launch_encode: division by mbw and by tiles_per_frame
    long long frame_stride;       // batches: bytes between consecutive frames' pixel arrays
    unsigned long long* frame_off;  // batches, optional: n_frames+1 payload offsets (two-pass mode)
    int sm_count;                 // filled by launch_encode
    int compact_cap;              // bytes of a tile that may be compacted in shared memory (0 = the built-in limit)
    int n_tiles;                  // filled by launch_encode: tiles of ENC_TILE_MB macroblocks covering the range
    uint8_t* out;                 // payload bytes for this range
    unsigned long long out_cap;
    unsigned long long* out_len;  // payload length, written by the last tile
    unsigned long long* out_len_user;   // two-pass mode, optional: a second place to write it
    // single-pass mode (stage == nullptr): decoupled look-back
    unsigned long long* desc;     // one look-back descriptor per tile, zeroed before launch
    unsigned int* ticket;         // zeroed before launch
    // two-pass mode (stage != nullptr): tiles land in a staging buffer in completion order, then
    // enc_scan_tiles + enc_gather_tiles move them to their payload offsets
    uint8_t* stage;               // 16-byte aligned
    unsigned long long stage_cap_units;   // capacity in 16-byte units
    unsigned int stage_slot_units;        // non-zero: tile t is staged at unit t * stage_slot_units (the buffer holds every tile's worst case)
    unsigned int* stage_cursor;   // zero before the first launch; enc_scan_tiles leaves it at zero again
    unsigned int* bucket;         // optional, ENC_BUCKETS words, zero before the first launch: byte counts per 64 tiles (left at zero again)
    unsigned int* gather_done;    // with `bucket`: one word, zero before the first launch
    uint2* tile_info;             // per tile: (staging offset in units, bytes)
    unsigned long long* tile_off; // per tile: payload offset
    const unsigned long long* chain_in;   // optional: payload offset to start at (sub-stripes of one frame, groups of a batch)
    unsigned long long* chain_out;        // optional: where the payload ended (a different word than chain_in)
    unsigned int* err;            // bit 0: out_cap too small
    uint32_t* chunk_off;          // optional: byte offset of every chunk in `out`
    int8_t* qcoef;                // optional: quantised coefficients, 64 per chunk, scan order
};
constexpr int ENC_BUCKETS = 1024;             // frames of up to 64 * ENC_BUCKETS tiles scan and gather in one kernel
size_t enc_smem_bytes();
cudaError_t launch_encode(const EncParams& P, int n_tiles, cudaStream_t stream);

// ---- decode
struct DecParams {
    const uint8_t* payload;
    unsigned long long payload_len;
    int width, height, mbw, mbh;
    int mb_row0, mb_row1;          // macroblock rows to reconstruct (parse covers everything before)
    uint8_t* dst;                  // address of output row 0 (bottom-up BGR) or of pixel (0,0) (RGB24)
    long long stride;              // bytes between output rows (BGR) — RGB24 uses 3*width
    int rgb_topdown;               // 0: bottom-up BGR rows (bmp.c:222-260); 1: packed top-down RGB24 (bmp.c:291-320)
};
struct DecWorkspace {
    // sized by dec_workspace_bytes(); all device memory
    uint8_t* base;
    size_t bytes;
};
size_t dec_workspace_bytes(unsigned long long payload_len, long long n_chunks);
// chunk_off_in: optional encoder side-info (skips boundary discovery when non-null)
// reuse_index: the workspace already holds this payload's chunk offsets and DCs (skip everything but reconstruct)
cudaError_t launch_decode(const DecParams& P, void* workspace, size_t workspace_bytes,
                          const uint32_t* chunk_off_in, bool reuse_index, unsigned int* err, cudaStream_t stream);

// ---- misc kernels (gpu_codec.cu)
cudaError_t launch_synth(uint8_t* rows, int width, int height, long long stride, int kind, uint32_t seed,
                         int frame, cudaStream_t stream);

}  // namespace awi
