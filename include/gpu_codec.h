/*
 * gpu_codec.h — C ABI of the B200 (sm_100a) AWI codec library, libgpu_codec.so.
 *
 * This is the boundary the reference's host code binds to.  The reference
 * (du207/image_compression) keeps its C host — bmp.c loading, the awi.c container,
 * the main.c CLI — and calls these entry points where its two macroblock loops were:
 *
 *     compress_to_awi   (awi.c:65-141)  loop body awi.c:94-128  -> gpu_codec_encode_bgr
 *     decompress_awi    (awi.c:153-230) loop body awi.c:184-217 -> gpu_codec_decode_to_bgr
 *                                                                  / gpu_codec_decode_to_rgb24
 *
 * Conventions follow the reference: plain pointers and sizes, `bool` results with
 * `true` = success (pixelrw.h:14, bitrw.h:16,23), one diagnostic line on stderr per
 * failure (awi.c:67,81,101,131), never abort().  CUDA failures surface as `false`.
 * Output buffers the library allocates are released with gpu_codec_free(), mirroring
 * the reference's "callee mallocs, caller frees" sinks (bmp.c:287, main.c:227).
 *
 * Results are bit-identical to the reference CPU codec built without FMA contraction
 * (its shipped -O0 build): same .awi payload bytes, same decoded pixels.
 *
 * Pixel layout ("bmp_rows"): exactly the pixel array of a 24-bit BMP file as the
 * reference reads it (bmp.c:35-36,80-81,111-113) — rows bottom-up, B,G,R bytes,
 * `row_stride` bytes apart (a BMP file uses ((24*width+31)/32)*4).
 *
 * Payload: the .awi file without its 16-byte header (the host writes that, awi.c:75-83):
 * for each 16x16 macroblock in raster order six byte-aligned chunks Y00 Y01 Y10 Y11 Cb Cr.
 *
 * There is no CPU fallback: every call fails (returns false) when no sm_100-class GPU
 * or no initialised context is available.
 */
#ifndef GPU_CODEC_H
#define GPU_CODEC_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ lifecycle */

/* Number of CUDA devices visible to this process (0 when there is none). */
int gpu_codec_device_count(void);

/* Create one context (stream, scratch, staging) on each of devices 0..n_gpus-1.
 * Large frames are split across them by macroblock-row stripes, batches by whole
 * frame (replaces nothing in the reference, which is single-threaded; the reference's
 * lazy cosine-table initialisation, dct.c:11-23, happens here instead). */
bool gpu_codec_init(int n_gpus);

/* Same, on an explicit device list (one process per GPU launches pass their own id).
 * Calling init again with the list already in use is a no-op that returns true. */
bool gpu_codec_init_devices(const int* device_ids, int n);

void gpu_codec_shutdown(void);

/* Text of the most recent failure on the calling thread's last call ("" if none). */
const char* gpu_codec_last_error(void);

/* Upper bound of the payload size for a width x height frame: 121 bytes per chunk
 * (12 + 63*(8+7) + 8 bits, rle.c:52-110 with |coefficient| <= 95). */
size_t gpu_codec_max_payload_bytes(int width, int height);

/* Page-locked host memory for the fastest transfers; optional: any host pointer works, ordinary (pageable) memory
 * is staged through page-locked slots inside the library. */
void* gpu_codec_host_alloc(size_t bytes);
void gpu_codec_host_free(void* p);

/* Releases buffers returned through `payload` / `payloads` below. */
void gpu_codec_free(void* p);

/* ------------------------------------------------------------------ host-buffer codec */

/* Replaces the macroblock loop of compress_to_awi (awi.c:85-128): colour transform
 * (color.c:14-36), 4:2:0 pick (color.c:69-86), DCT + quantise + scan (dct.c:26-53,
 * 109-126,159-169), run-length coding (rle.c:52-110), bit packing (bitrw.c:9-43) and
 * the six DC predictors (awi.c:91-92).  *payload is allocated by the library (page-locked); release it with
 * gpu_codec_free. */
bool gpu_codec_encode_bgr(const uint8_t* bmp_rows, int width, int height, int row_stride,
                          uint8_t** payload, size_t* payload_len);

/* Image-parallel batch of equally sized frames (whole frames per GPU, predictors reset
 * per frame — n independent compress_to_awi calls).  payloads[i] is allocated by the library, one gpu_codec_free each. */
bool gpu_codec_encode_batch_bgr(const uint8_t* const* frames, int n_frames, int width, int height,
                                int row_stride, uint8_t** payloads, size_t* payload_lens);

/* Replaces the macroblock loop of decompress_awi (awi.c:175-217) with the BMP sink's
 * layout (bmp.c:222-260): read_and_rle_decode_chunk (rle.c:113-170), inverse scan and
 * dequantise (dct.c:128-145,171-181), IDCT (dct.c:55-81), 4:2:0 replicate (color.c:88-105),
 * inverse colour (color.c:45-67), crop.  Row padding bytes are left untouched.  The payload must hold all
 * 6 x macroblocks chunks; bytes after the last chunk are ignored, as the reference does. */
bool gpu_codec_decode_to_bgr(const uint8_t* payload, size_t len, int width, int height,
                             uint8_t* bmp_rows, int row_stride);

/* Same with the RGB24 sink's layout (bmp.c:291-320; `show`, main.c:162-181): packed
 * top-down R,G,B, pitch 3*width. */
bool gpu_codec_decode_to_rgb24(const uint8_t* payload, size_t len, int width, int height,
                               uint8_t* rgb_topdown);

/* ------------------------------------------------------------------ device-resident codec
 * All pointers below are device pointers on the context's device `ctx` (index into the
 * list given to init).  Calls are asynchronous on `stream` (a cudaStream_t; NULL is the
 * CUDA default stream, as everywhere in CUDA) and must be stream-ordered per context: two
 * encodes (or two decodes) of one context in flight on different streams race on its
 * scratch. */

/* Encode macroblock rows [mb_row0, mb_row1) of a width x height frame.  d_rows holds
 * the frame's pixel rows window_y0 <= y < window_y1 (top-down numbering) in bottom-up
 * order, i.e. its first byte is the first byte of row window_y1-1.  The window must
 * cover rows 16*mb_row0-16 (clamped to 0) .. min(height,16*mb_row1): the extra
 * macroblock row above supplies the DC predictors at a stripe seam.
 * d_payload_len (8 bytes) receives the stripe's payload length; d_chunk_off (optional,
 * one uint32 per chunk) each chunk's byte offset; d_qcoef (optional, 64 int8 per chunk)
 * the quantised coefficients in chunk order (test/diagnostic hook). */
bool gpu_codec_encode_bgr_dev(int ctx, const uint8_t* d_rows, int width, int height, long long row_stride,
                              int window_y0, int window_y1, int mb_row0, int mb_row1,
                              uint8_t* d_payload, size_t payload_cap, uint64_t* d_payload_len,
                              uint32_t* d_chunk_off, int8_t* d_qcoef, void* stream);

/* Image-parallel batch on one device: n_frames equally sized frames, frame_stride bytes apart, each
 * encoded as its own stream (DC predictors restart per frame, awi.c:91-92) by ONE set of kernel
 * launches.  The payloads are concatenated in d_payload in frame order; d_frame_off receives
 * n_frames+1 offsets (frame i occupies [off[i], off[i+1])). */
bool gpu_codec_encode_batch_bgr_dev(int ctx, const uint8_t* d_frames, int n_frames, long long frame_stride,
                                    int width, int height, long long row_stride,
                                    uint8_t* d_payload, size_t payload_cap, uint64_t* d_frame_off, void* stream);

/* Whole-frame encode of device-resident rows that allocates nothing (no counterpart in
 * the reference; it exists for callers of this library that must not allocate device
 * memory while encoding: engines that own all device memory through their own pool,
 * loops that cannot afford the device-wide stall of cudaMalloc/cudaFree, code running
 * while another stream is under graph capture).  Same kernel and same bytes as
 * gpu_codec_encode_bgr_dev on the full frame, but the kernel's scratch (tile ticket +
 * one 8-byte descriptor per tile of 32 macroblocks) is a fixed 512 KB + 256 B array
 * inside the library, so:
 *   - limit: at most 65536 tiles = 2,097,152 macroblocks per call (e.g. 16384 x 16384 is
 *     1,048,576); larger frames are refused with an error, nothing is launched;
 *   - no context and no gpu_codec_init are needed; the frame is encoded on the CURRENT
 *     device, all pointers are device pointers there;
 *   - the first call on each device looks up the array's device address; make that
 *     first call before starting any stream capture;
 *   - all calls on one device share the array: two calls whose kernels overlap on the
 *     GPU (different streams, no ordering) corrupt each other; order them;
 *   - d_payload_len (8 bytes) always receives the length the payload needs; d_status (4 bytes)
 *     is 0 on success, or has bit 0 set when payload_cap was smaller than that length.  In that
 *     case the tiles that still fitted have been written, so the contents of d_payload are
 *     unspecified: retry with a buffer of at least *d_payload_len bytes
 *     (gpu_codec_max_payload_bytes always suffices).  Both words are written on `stream`; read
 *     them after synchronising it.
 * The sizes above are for the library as shipped (32 macroblocks per tile); the source asserts it. */
bool gpu_codec_encode_bgr_dev_noalloc(const uint8_t* d_rows, int width, int height, long long row_stride,
                                      uint8_t* d_payload, size_t payload_cap, uint64_t* d_payload_len,
                                      uint32_t* d_status, void* stream);

/* Decode a whole payload's chunk structure, reconstruct macroblock rows
 * [mb_row0, mb_row1) ("replicated parse, sharded reconstruct").  d_dst is the address
 * of bottom-up row 0 (BGR sink) or of pixel (0,0) (RGB24 sink, rgb_topdown != 0) of the
 * FULL frame; only the requested rows are written.  d_chunk_off (optional) is the
 * encoder's side-info and skips chunk-boundary discovery. */
bool gpu_codec_decode_dev(int ctx, const uint8_t* d_payload, size_t len, int width, int height,
                          int mb_row0, int mb_row1, uint8_t* d_dst, long long row_stride,
                          int rgb_topdown, const uint32_t* d_chunk_off, void* stream);

/* Wait for `stream` on context `ctx` and report whether any call since the last check
 * flagged an error (payload_cap too small, malformed stream, CUDA error). */
bool gpu_codec_sync_check(int ctx, void* stream);

/* ------------------------------------------------------------------ synthetic frames + self-tests
 * Everything below this line serves the repository's own tests and bench (synthetic inputs, exhaustive colour
 * check, a knob that forces a rarely taken write path).  It is NOT part of the stable drop-in surface: a host
 * that binds the codec needs none of it, and names or behaviour may change between versions. */

/* Fill d_rows (bottom-up BGR, row_stride apart) with an integer-only synthetic frame;
 * kind 0 = noise, 1 = photo-like, 2 = flat 144 (identical to image_compression_b200/synth.py). */
bool gpu_codec_synth_bgr_dev(int ctx, uint8_t* d_rows, int width, int height, long long row_stride,
                             int kind, uint32_t seed, int frame, void* stream);

/* Colour transform of n packed triples on the device, for exhaustive parity tests:
 * inverse == 0: (R,G,B) -> (Y,Cb,Cr) (color.c:31-33); else (Y,Cb,Cr) -> (R,G,B) (color.c:61-63). */
bool gpu_codec_selftest_colour_dev(int ctx, const uint8_t* d_in, size_t n, int inverse, uint8_t* d_out, void* stream);

/* Test hook: pretend the encoder's per-tile shared-memory buffer holds only `bytes` (0 restores the real
 * 12 KB), which forces the chunk-by-chunk write path that otherwise needs > 64 bytes per chunk on average
 * across a whole tile.  Results are unaffected.  No counterpart in the reference.  Process-wide and not
 * synchronised: set it only while no encode call is running (tests only). */
void gpu_codec_selftest_tile_buffer(int bytes);

#ifdef __cplusplus
}
#endif
#endif /* GPU_CODEC_H */
