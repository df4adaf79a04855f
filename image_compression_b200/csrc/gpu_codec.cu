// gpu_codec.cu — the C ABI declared in include/gpu_codec.h: contexts, scratch and staging
// management, stripe / batch partitioning over GPUs and host-side stitching.  No codec
// arithmetic lives here (see awi_encode.cu / awi_decode.cu) and there is no CPU path:
// without a usable CUDA device every entry point reports failure.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <condition_variable>
#if defined(__x86_64__)
#include <emmintrin.h>
#endif
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/gpu_codec.h"
#include "awi_kernels.h"
#include "codec_math.cuh"

namespace {

thread_local std::string g_last_error;

bool fail(const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_last_error = buf;
    fprintf(stderr, "gpu_codec: %s\n", buf);      // reference convention: one stderr line, return false
    return false;
}

#define CU(call)                                                                                  \
    do {                                                                                          \
        cudaError_t e__ = (call);                                                                 \
        if (e__ != cudaSuccess) return fail("%s failed: %s", #call, cudaGetErrorString(e__));     \
    } while (0)

constexpr int MAX_SUB = 32;               // sub-stripes a frame is cut into to overlap PCIe copies with the kernels
constexpr size_t SUB_BYTES = 48u << 20;   // aim for ~48 MB of pixels per sub-stripe
constexpr size_t SUB_BYTES_MIN = 4u << 20;
constexpr int RING_SLOTS = 3;              // page-locked staging slots per context for uploads from pageable memory
constexpr size_t ENC_WS_HEADER = 256 + 4 * awi::ENC_BUCKETS;   // [0,8) payload length, [8,12) tile ticket, [16,20) staging cursor, [24,28) gather count,
                                                             // [256, ...) byte counts per 64 tiles

struct Buf {
    uint8_t* p = nullptr;
    size_t bytes = 0;
};

struct Ctx {
    int dev = -1;
    cudaStream_t stream = nullptr;            // compute (and small copies)
    cudaStream_t copy_stream = nullptr;       // bulk H2D / D2H overlapped with compute
    cudaStream_t back_stream = nullptr;       // payload D2H while later sub-stripes are still uploading
    cudaEvent_t ev[MAX_SUB] = { nullptr };
    cudaEvent_t ev_done[MAX_SUB] = { nullptr };
    unsigned long long* h_chain = nullptr;    // pinned: payload length after each sub-stripe
    Buf enc_ws, enc_stage, dec_ws, in, out, frame_off;
    unsigned int* d_err = nullptr;
    unsigned long long* d_run = nullptr;      // two words: running payload offset of a frame encoded in sub-stripes, read from one
                                              // and written to the other by each launch (alternating)
    uint64_t* h_len = nullptr;                // pinned
    // uploads from pageable memory go through these page-locked slots (see upload())
    uint8_t* ring[RING_SLOTS] = { nullptr };
    size_t ring_bytes = 0;
    cudaEvent_t ring_ev[RING_SLOTS] = { nullptr };
    bool ring_busy[RING_SLOTS] = { false };
    int ring_next = 0;
    std::mutex mu;
};

int g_selftest_compact_cap = 0;           // see gpu_codec_selftest_tile_buffer
std::vector<Ctx*> g_ctx;
std::mutex g_mu;

bool ensure(Buf& b, size_t bytes) {
    if (b.bytes >= bytes) return true;
    if (b.p) CU(cudaFree(b.p));
    b.p = nullptr;
    b.bytes = 0;
    size_t want = bytes + bytes / 8 + 4096;
    CU(cudaMalloc(&b.p, want));
    b.bytes = want;
    return true;
}

Ctx* get_ctx(int i) {
    if (i < 0 || i >= (int) g_ctx.size()) {
        fail("context %d not initialised (call gpu_codec_init first)", i);
        return nullptr;
    }
    return g_ctx[i];
}

void destroy_ctx(Ctx* c) {
    if (!c) return;
    cudaSetDevice(c->dev);
    if (c->stream) cudaStreamDestroy(c->stream);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->back_stream) cudaStreamDestroy(c->back_stream);
    for (cudaEvent_t e : c->ev) if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : c->ev_done) if (e) cudaEventDestroy(e);
    if (c->h_chain) cudaFreeHost(c->h_chain);
    if (c->d_run) cudaFree(c->d_run);
    for (Buf* b : { &c->enc_ws, &c->enc_stage, &c->dec_ws, &c->in, &c->out, &c->frame_off })
        if (b->p) cudaFree(b->p);
    if (c->d_err) cudaFree(c->d_err);
    if (c->h_len) cudaFreeHost(c->h_len);
    for (int i = 0; i < RING_SLOTS; i++) {
        if (c->ring[i]) cudaFreeHost(c->ring[i]);
        if (c->ring_ev[i]) cudaEventDestroy(c->ring_ev[i]);
    }
    delete c;
}

bool make_events(Ctx* c) {
    for (int i = 0; i < MAX_SUB; i++)
        if (cudaEventCreateWithFlags(&c->ev[i], cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&c->ev_done[i], cudaEventDisableTiming) != cudaSuccess)
            return false;
    return cudaStreamCreateWithFlags(&c->back_stream, cudaStreamNonBlocking) == cudaSuccess &&
           cudaMallocHost(&c->h_chain, MAX_SUB * 8) == cudaSuccess;
}

// Result buffers handed to the caller (`*payload`) come from a small pool of page-locked
// blocks so the device->host copy runs at full PCIe speed and no page is pinned per call.
// gpu_codec_free() puts a block back; anything not from the pool goes to free().
struct PinnedBlock { void* p; size_t cap; bool in_use; };
std::vector<PinnedBlock> g_pool;
std::mutex g_pool_mu;

void* pool_get(size_t bytes) {
    std::lock_guard<std::mutex> lk(g_pool_mu);
    if (bytes == 0) bytes = 1;
    for (PinnedBlock& b : g_pool)
        if (!b.in_use && b.cap >= bytes && b.cap <= 4 * bytes + (1u << 20)) { b.in_use = true; return b.p; }
    void* p = nullptr;
    const size_t cap = bytes + bytes / 4 + 4096;
    if (cudaMallocHost(&p, cap) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    g_pool.push_back({ p, cap, true });
    return p;
}

bool pool_put(void* p) {           // true if p was a pool block
    std::lock_guard<std::mutex> lk(g_pool_mu);
    size_t idle = 0;
    for (PinnedBlock& b : g_pool) if (!b.in_use) idle += b.cap;
    for (size_t i = 0; i < g_pool.size(); i++)
        if (g_pool[i].p == p) {
            if (idle > (1ull << 30)) {             // do not hoard more than 1 GB of idle pinned memory
                cudaFreeHost(p);
                g_pool.erase(g_pool.begin() + (long) i);
            } else {
                g_pool[i].in_use = false;
            }
            return true;
        }
    return false;
}

void pool_trim() {
    std::lock_guard<std::mutex> lk(g_pool_mu);
    for (size_t i = 0; i < g_pool.size();)
        if (!g_pool[i].in_use) { cudaFreeHost(g_pool[i].p); g_pool.erase(g_pool.begin() + (long) i); } else i++;
}

inline int align_of(const void* p, long long stride) {
    uintptr_t a = reinterpret_cast<uintptr_t>(p);
    if (a % 16 == 0 && stride % 16 == 0) return 16;
    if (a % 4 == 0 && stride % 4 == 0) return 4;
    return 1;
}

// -------------------------------------------------------------- uploads from pageable memory
//
// The reference reads its BMP into ordinary malloc'd memory (bmp.c:38, :125), and a drop-in caller will hand us
// exactly that.  cudaMemcpyAsync from pageable memory is staged by the driver on one thread (5-10 GB/s measured
// here, against 55 GB/s from page-locked memory), so such sources are staged here instead: a few page-locked
// slots per context, filled by a small pool of copy threads while the previous slot is on the wire.
// Large copies between pageable memory and the ring bypass the cache on the store side (the bytes are not read
// again by this core: the copy engine or the caller takes them), which saves the read-for-ownership traffic.
static void stream_copy(uint8_t* d, const uint8_t* s, size_t n) {
#if defined(__x86_64__)
    if (n >= 4096) {
        size_t head = (16 - (reinterpret_cast<uintptr_t>(d) & 15)) & 15;
        memcpy(d, s, head);
        d += head; s += head; n -= head;
        const size_t blocks = n / 64;
        for (size_t i = 0; i < blocks; i++) {
            const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i*>(s)), b = _mm_loadu_si128(reinterpret_cast<const __m128i*>(s + 16));
            const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i*>(s + 32)), e = _mm_loadu_si128(reinterpret_cast<const __m128i*>(s + 48));
            _mm_stream_si128(reinterpret_cast<__m128i*>(d), a);
            _mm_stream_si128(reinterpret_cast<__m128i*>(d + 16), b);
            _mm_stream_si128(reinterpret_cast<__m128i*>(d + 32), c);
            _mm_stream_si128(reinterpret_cast<__m128i*>(d + 48), e);
            d += 64; s += 64;
        }
        _mm_sfence();
        n -= blocks * 64;
    }
#endif
    memcpy(d, s, n);
}

class CopyPool {
public:
    static CopyPool& get() { static CopyPool* p = new CopyPool; return *p; }      // never destroyed: its threads sleep until exit
    // dst/src are split into one piece per worker; returns when every piece is copied
    void copy(uint8_t* dst, const uint8_t* src, size_t n) {
        const size_t nw = workers_.size();
        if (nw == 0 || n < (4u << 20)) { memcpy(dst, src, n); return; }
        std::unique_lock<std::mutex> call(call_mu_);          // one parallel copy at a time
        {
            std::lock_guard<std::mutex> lk(mu_);
            const size_t piece = ((n + nw - 1) / nw + 4095) & ~(size_t) 4095;
            for (size_t i = 0; i < nw; i++) {
                const size_t a = i * piece < n ? i * piece : n, b = a + piece < n ? a + piece : n;
                jobs_[i] = { dst + a, src + a, b - a };
            }
            pending_ = (int) nw;
            generation_++;
        }
        cv_.notify_all();
        std::unique_lock<std::mutex> lk(mu_);
        done_.wait(lk, [&] { return pending_ == 0; });
    }
private:
    struct Job { uint8_t* d; const uint8_t* s; size_t n; };
    CopyPool() {
        unsigned hw = std::thread::hardware_concurrency();
        unsigned nw = hw >= 16 ? 8 : (hw >= 4 ? hw / 2 : 0);
        jobs_.resize(nw);
        for (unsigned i = 0; i < nw; i++) workers_.emplace_back([this, i] { run(i); });
        for (auto& t : workers_) t.detach();                  // they sleep on cv_ for the life of the process
    }
    void run(unsigned i) {
        unsigned long long seen = 0;
        while (true) {
            Job j;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return generation_ != seen; });
                seen = generation_;
                j = jobs_[i];
            }
            if (j.n) stream_copy(j.d, j.s, j.n);
            {
                std::lock_guard<std::mutex> lk(mu_);
                if (--pending_ == 0) done_.notify_all();
            }
        }
    }
    std::vector<std::thread> workers_;
    std::vector<Job> jobs_;
    std::mutex mu_, call_mu_;
    std::condition_variable cv_, done_;
    unsigned long long generation_ = 0;
    int pending_ = 0;
};

bool is_pageable(const void* p) {
    static const bool off = getenv("GPU_CODEC_NO_STAGING") != nullptr;      // measurement knob: leave pageable memory to the driver
    if (off) return false;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return true; }
    return a.type == cudaMemoryTypeUnregistered;
}

// Host -> device on the context's copy stream.  Page-locked sources go straight to the copy engine; pageable ones
// pass through the ring (at most `ring_bytes` per slot: larger requests are cut).
bool upload(Ctx* c, uint8_t* d_dst, const uint8_t* h_src, size_t n, bool pageable) {
    if (!pageable) {
        CU(cudaMemcpyAsync(d_dst, h_src, n, cudaMemcpyHostToDevice, c->copy_stream));
        return true;
    }
    const size_t want = 32u << 20;
    if (c->ring_bytes == 0) {
        for (int i = 0; i < RING_SLOTS; i++) {
            CU(cudaMallocHost(&c->ring[i], want));
            CU(cudaEventCreateWithFlags(&c->ring_ev[i], cudaEventDisableTiming));
        }
        c->ring_bytes = want;
    }
    for (size_t at = 0; at < n; at += c->ring_bytes) {
        const size_t m = n - at < c->ring_bytes ? n - at : c->ring_bytes;
        const int slot = c->ring_next;
        c->ring_next = (slot + 1) % RING_SLOTS;
        if (c->ring_busy[slot]) CU(cudaEventSynchronize(c->ring_ev[slot]));      // its previous contents are on the device
        CopyPool::get().copy(c->ring[slot], h_src + at, m);
        CU(cudaMemcpyAsync(d_dst + at, c->ring[slot], m, cudaMemcpyHostToDevice, c->copy_stream));
        CU(cudaEventRecord(c->ring_ev[slot], c->copy_stream));
        c->ring_busy[slot] = true;
    }
    return true;
}

// Device -> pageable host memory through the same ring: the copy engine fills a slot while the copy threads
// empty the one before it.  `pend` remembers slots whose bytes still have to be moved to their destination.
struct PendingOut { int slot; uint8_t* h_dst; size_t n; };
bool drain_one(Ctx* c, std::vector<PendingOut>& pend) {
    if (pend.empty()) return true;
    const PendingOut p = pend.front();
    pend.erase(pend.begin());
    CU(cudaEventSynchronize(c->ring_ev[p.slot]));
    CopyPool::get().copy(p.h_dst, c->ring[p.slot], p.n);
    c->ring_busy[p.slot] = false;
    return true;
}
bool download(Ctx* c, uint8_t* h_dst, const uint8_t* d_src, size_t n, std::vector<PendingOut>& pend) {
    const size_t want = 32u << 20;
    if (c->ring_bytes == 0) {
        for (int i = 0; i < RING_SLOTS; i++) {
            CU(cudaMallocHost(&c->ring[i], want));
            CU(cudaEventCreateWithFlags(&c->ring_ev[i], cudaEventDisableTiming));
        }
        c->ring_bytes = want;
    }
    const size_t piece = 8u << 20;                             // small pieces: only the last one's copy-out is not overlapped
    for (size_t at = 0; at < n; at += piece) {
        const size_t m = n - at < piece ? n - at : piece;
        const int slot = c->ring_next;
        c->ring_next = (slot + 1) % RING_SLOTS;
        while (!pend.empty() && c->ring_busy[slot]) {          // the slot still holds undelivered bytes (oldest first)
            bool mine = false;
            for (const PendingOut& p : pend) mine = mine || p.slot == slot;
            if (!mine) break;
            if (!drain_one(c, pend)) return false;
        }
        if (c->ring_busy[slot]) CU(cudaEventSynchronize(c->ring_ev[slot]));      // (an earlier upload from it)
        CU(cudaMemcpyAsync(c->ring[slot], d_src + at, m, cudaMemcpyDeviceToHost, c->copy_stream));
        CU(cudaEventRecord(c->ring_ev[slot], c->copy_stream));
        c->ring_busy[slot] = true;
        pend.push_back({ slot, h_dst + at, m });
        while (pend.size() > 1)                                // deliver the piece before while this one is on the wire
            if (!drain_one(c, pend)) return false;
    }
    return true;
}

// -------------------------------------------------------------- small kernels

__global__ void synth_kernel(uint8_t* rows, int width, int height, long long stride, int kind, uint32_t seed, int frame) {
    // image_compression_b200/synth.py restated; one thread per byte of the pixel array
    long long n = (long long) width * height * 3;
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        long long pix = i / 3;
        int c = (int) (i - pix * 3);             // 0,1,2 = R,G,B
        int y = (int) (pix / width), x = (int) (pix - (long long) y * width);
        int v;
        if (kind == 2) {
            v = 144;
        } else {
            uint32_t s = seed ^ ((uint32_t) frame * 0x9E3779B9u);
            uint32_t h = (uint32_t) i * 0x9E3779B1u + s;
            h ^= h >> 16; h *= 0x85EBCA6Bu; h ^= h >> 13; h *= 0xC2B2AE35u; h ^= h >> 16;
            if (kind == 0) {
                v = (int) (h >> 24);
            } else {
                int t1 = 3 * x + 37 * c + 5 * frame, t2 = 2 * y + 91 * c;
                int tri1 = abs((t1 & 511) - 256) - 128, tri2 = abs((t2 & 511) - 256) - 128;
                v = 128 + ((tri1 + tri2) >> 1) + ((int) (h >> 27) - 16);
                v = v < 0 ? 0 : (v > 255 ? 255 : v);
            }
        }
        rows[(long long) (height - 1 - y) * stride + 3ll * x + (2 - c)] = (uint8_t) v;
    }
}

__global__ void pad_zero_kernel(uint8_t* rows, int width, int height, long long stride) {
    int pad = (int) (stride - 3ll * width);
    long long n = (long long) pad * height;
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x)
        rows[(i / pad) * stride + 3ll * width + (i % pad)] = 0;
}

__global__ void colour_kernel(const uint8_t* in, size_t n, int inverse, uint8_t* out) {
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
        int a = in[3 * i], b = in[3 * i + 1], c = in[3 * i + 2];
        if (!inverse) {
            out[3 * i] = (uint8_t) awi::luma_of(a, b, c);
            out[3 * i + 1] = (uint8_t) awi::cb_of(a, b, c);
            out[3 * i + 2] = (uint8_t) awi::cr_of(a, b, c);
        } else {
            out[3 * i] = (uint8_t) awi::red_of(a, c);
            out[3 * i + 1] = (uint8_t) awi::green_of(a, b, c);
            out[3 * i + 2] = (uint8_t) awi::blue_of(a, b);
        }
    }
}

// -------------------------------------------------------------- device-resident bodies

bool encode_dev_impl(Ctx* c, const uint8_t* d_rows, int width, int height, long long stride, int wy0, int wy1,
                     int r0, int r1, uint8_t* d_payload, size_t cap, uint64_t* d_len, uint32_t* d_chunk_off,
                     int8_t* d_qcoef, cudaStream_t st, unsigned long long* d_chain = nullptr, int chain_slot = 0, int n_frames = 1,
                     long long frame_stride = 0, unsigned long long* d_frame_off = nullptr) {
    if (width <= 0 || height <= 0) return fail("encode: bad geometry %dx%d", width, height);
    if (stride < 3ll * width) return fail("encode: row_stride %lld < 3*width", stride);
    const int mbw = (width + 15) / 16, mbh = (height + 15) / 16;
    if (r0 < 0 || r1 > mbh || r0 >= r1) return fail("encode: bad macroblock row range [%d,%d) of %d", r0, r1, mbh);
    if ((long long) mbw * mbh > (1ll << 30)) return fail("encode: frame too large");
    const int need_lo = r0 > 0 ? 16 * r0 - 16 : 0, need_hi = 16 * r1 < height ? 16 * r1 : height;
    if (wy0 > need_lo || wy1 < need_hi || wy1 > height || wy0 < 0)
        return fail("encode: pixel window [%d,%d) does not cover rows [%d,%d)", wy0, wy1, need_lo, need_hi);
    if (!d_rows || !d_payload) return fail("encode: null device pointer");
    if (d_chunk_off && cap >= (1ull << 32)) return fail("encode: chunk offsets are 32-bit; payload_cap must stay below 4 GiB when they are requested");

    const long long n_mb = (long long) (r1 - r0) * mbw;
    const int tiles_per_frame = (int) ((n_mb + awi::ENC_TILE_MB - 1) / awi::ENC_TILE_MB);
    if (n_frames < 1 || (long long) tiles_per_frame * n_frames > (1ll << 30)) return fail("encode: bad frame count %d", n_frames);
    const int n_tiles = tiles_per_frame * n_frames;
    CU(cudaSetDevice(c->dev));
    // two-pass encode: tiles go to a staging buffer in completion order (no tile ever waits for
    // another), then a scan + gather place them; scratch = header + (where,bytes) + offset per tile
    const size_t ws_bytes = ENC_WS_HEADER + (size_t) n_tiles * 16;
    // a payload capacity that covers the worst case (every chunk 121 bytes) lets every tile have its own staging slot;
    // a tighter capacity (the caller knows its content) keeps the staging buffer as small as the payload
    constexpr size_t SLOT_UNITS = ((size_t) awi::ENC_TILE_MB * 6 * 121 + 15) / 16;
    static const bool slots_allowed = getenv("GPU_CODEC_NO_STAGE_SLOTS") == nullptr;      // (measurement knob)
    const bool slots = slots_allowed && cap >= (size_t) n_mb * n_frames * 6 * 121 && (size_t) n_tiles * SLOT_UNITS < (1ull << 32);
    const size_t stage_bytes = slots ? (size_t) n_tiles * SLOT_UNITS * 16 + 256 : cap + (size_t) n_tiles * 16 + 256;
    const size_t ws_before = c->enc_ws.bytes;
    if (!ensure(c->enc_ws, ws_bytes) || !ensure(c->enc_stage, stage_bytes)) return false;
    // the staging cursor must be zero when a launch starts: once here for a fresh buffer, afterwards the scan
    // kernel of every launch leaves it at zero (one operation less per call)
    if (c->enc_ws.bytes != ws_before) CU(cudaMemsetAsync(c->enc_ws.p, 0, ENC_WS_HEADER, st));      // ordered with the kernels that follow on `st`

    awi::EncParams P{};
    // address of bottom-up row 0 of the full frame, had the whole frame been resident
    P.rows = d_rows - (long long) (height - wy1) * stride;
    P.stride = stride;
    P.width = width;
    P.height = height;
    P.align = align_of(d_rows, stride);
    if (n_frames > 1 && frame_stride % P.align != 0)         // every frame's rows must be as aligned as the first one's
        P.align = frame_stride % 4 == 0 && P.align >= 4 ? 4 : 1;
    P.mbw = mbw;
    P.mb0 = r0 * mbw;
    P.mb1 = r1 * mbw;
    P.out = d_payload;
    P.out_cap = cap;
    P.out_len = reinterpret_cast<unsigned long long*>(c->enc_ws.p);
    P.ticket = reinterpret_cast<unsigned int*>(c->enc_ws.p + 8);
    P.desc = nullptr;
    P.stage = c->enc_stage.p;
    P.stage_cap_units = (stage_bytes - 256) / 16;
    P.stage_slot_units = slots ? (unsigned int) SLOT_UNITS : 0u;
    P.stage_cursor = reinterpret_cast<unsigned int*>(c->enc_ws.p + 16);
    P.gather_done = reinterpret_cast<unsigned int*>(c->enc_ws.p + 24);
    P.bucket = reinterpret_cast<unsigned int*>(c->enc_ws.p + 256);
    P.tile_info = reinterpret_cast<uint2*>(c->enc_ws.p + ENC_WS_HEADER);
    P.tile_off = reinterpret_cast<unsigned long long*>(c->enc_ws.p + ENC_WS_HEADER + (size_t) n_tiles * 8);
    P.chain_in = d_chain ? d_chain + (chain_slot & 1) : nullptr;
    P.chain_out = d_chain ? d_chain + ((chain_slot & 1) ^ 1) : nullptr;
    P.compact_cap = g_selftest_compact_cap;
    P.tiles_per_frame = tiles_per_frame;
    P.frame_stride = frame_stride;
    P.frame_off = d_frame_off;
    P.err = c->d_err;
    P.chunk_off = d_chunk_off;
    P.qcoef = d_qcoef;
    P.out_len_user = reinterpret_cast<unsigned long long*>(d_len);
    CU(awi::launch_encode(P, n_tiles, st));
    return true;
}

bool decode_dev_impl(Ctx* c, const uint8_t* d_payload, size_t len, int width, int height, int r0, int r1,
                     uint8_t* d_dst, long long stride, int rgb, const uint32_t* d_chunk_off, cudaStream_t st,
                     bool reuse_index = false) {
    if (width <= 0 || height <= 0) return fail("decode: bad geometry %dx%d", width, height);
    const int mbw = (width + 15) / 16, mbh = (height + 15) / 16;
    if (r0 < 0 || r1 > mbh || r0 > r1) return fail("decode: bad macroblock row range [%d,%d) of %d", r0, r1, mbh);
    if (!rgb && stride < 3ll * width) return fail("decode: row_stride %lld < 3*width", stride);
    if (len >= (1ull << 32) - 256) return fail("decode: payload of %zu bytes is too large", len);
    const long long n_chunks = (long long) mbw * mbh * 6;
    if (len < (size_t) n_chunks * 3) return fail("decode: payload of %zu bytes is shorter than %lld chunks", len, n_chunks);
    if (!d_payload || !d_dst) return fail("decode: null device pointer");
    CU(cudaSetDevice(c->dev));
    const size_t ws = awi::dec_workspace_bytes(len, n_chunks);
    if (!ensure(c->dec_ws, ws)) return false;
    awi::DecParams P{};
    P.payload = d_payload;
    P.payload_len = len;
    P.width = width;
    P.height = height;
    P.mbw = mbw;
    P.mbh = mbh;
    P.mb_row0 = r0;
    P.mb_row1 = r1;
    P.dst = d_dst;
    P.stride = stride;
    P.rgb_topdown = rgb ? 1 : 0;
    CU(awi::launch_decode(P, c->dec_ws.p, c->dec_ws.bytes, d_chunk_off, reuse_index, c->d_err, st));
    return true;
}

bool check_err(Ctx* c, cudaStream_t st, const char* what) {
    CU(cudaSetDevice(c->dev));
    // the flags ride home behind whatever is queued on the stream: one wait instead of a wait and a blocking copy
    volatile unsigned int* h_err = reinterpret_cast<volatile unsigned int*>(c->h_len + 4);
    CU(cudaMemcpyAsync(const_cast<unsigned int*>(h_err), c->d_err, 4, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const unsigned int e = *h_err;
    if (e) {
        CU(cudaMemsetAsync(c->d_err, 0, 4, st));      // on the stream the flags came back on: cannot wipe a later call's flag
        CU(cudaStreamSynchronize(st));
        return fail("%s: device reported%s%s%s", what, (e & 1) ? " [payload buffer too small]" : "",
                    (e & 2) ? " [malformed stream: chunk boundaries]" : "", (e & 4) ? " [malformed stream: chunk body]" : "");
    }
    return true;
}

// Failure paths: nothing may still be reading the caller's pixels or writing into a block that is about to be
// handed back when a host entry point returns, so every stream of the context is drained first.
void drain(Ctx* c) {
    cudaSetDevice(c->dev);
    cudaStreamSynchronize(c->copy_stream);
    cudaStreamSynchronize(c->stream);
    cudaStreamSynchronize(c->back_stream);
    cudaGetLastError();
}

// rows of stripe g of G over mbh macroblock rows
inline void stripe_rows(int mbh, int g, int G, int& r0, int& r1) {
    r0 = (int) ((long long) mbh * g / G);
    r1 = (int) ((long long) mbh * (g + 1) / G);
}

}  // namespace

// ================================================================== C ABI
extern "C" {

int gpu_codec_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

bool gpu_codec_init_devices(const int* ids, int n) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_ctx.empty()) {                       // idempotent for the same device list
        bool same = ids && n == (int) g_ctx.size();
        for (int i = 0; same && i < n; i++) same = g_ctx[i]->dev == ids[i];
        if (same) return true;
        return fail("already initialised on a different device list (call gpu_codec_shutdown first)");
    }
    int have = gpu_codec_device_count();
    if (have == 0) return fail("no CUDA device available; this library has no CPU path");
    if (n <= 0 || !ids) return fail("init: need at least one device");
    for (int i = 0; i < n; i++) {
        if (ids[i] < 0 || ids[i] >= have) { return fail("init: device %d not present (%d visible)", ids[i], have); }
        cudaDeviceProp prop;
        CU(cudaGetDeviceProperties(&prop, ids[i]));
        if (prop.major != 10) return fail("init: device %d is sm_%d%d; this library is built for sm_100a only", ids[i], prop.major, prop.minor);
    }
    for (int i = 0; i < n; i++) {
        Ctx* c = new Ctx();
        c->dev = ids[i];
        g_ctx.push_back(c);
        if (cudaSetDevice(c->dev) != cudaSuccess || cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
            cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking) != cudaSuccess ||
            cudaMalloc(&c->d_run, 16) != cudaSuccess || !make_events(c) ||
            cudaMalloc(&c->d_err, 4) != cudaSuccess || cudaMemset(c->d_err, 0, 4) != cudaSuccess ||
            cudaMallocHost(&c->h_len, 64) != cudaSuccess) {
            cudaError_t e = cudaGetLastError();
            for (Ctx* k : g_ctx) destroy_ctx(k);
            g_ctx.clear();
            return fail("init: context creation on device %d failed: %s", ids[i], cudaGetErrorString(e));
        }
    }
    return true;
}

bool gpu_codec_init(int n_gpus) {
    if (n_gpus <= 0) return fail("init: n_gpus must be positive");
    std::vector<int> ids(n_gpus);
    for (int i = 0; i < n_gpus; i++) ids[i] = i;
    return gpu_codec_init_devices(ids.data(), n_gpus);
}

void gpu_codec_shutdown(void) {
    std::lock_guard<std::mutex> lk(g_mu);
    for (Ctx* c : g_ctx) destroy_ctx(c);
    g_ctx.clear();
    pool_trim();
}

const char* gpu_codec_last_error(void) { return g_last_error.c_str(); }

size_t gpu_codec_max_payload_bytes(int width, int height) {
    if (width <= 0 || height <= 0) return 0;
    return (size_t) ((width + 15) / 16) * (size_t) ((height + 15) / 16) * 6 * 121;
}

void* gpu_codec_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaMallocHost(&p, bytes) != cudaSuccess) { cudaGetLastError(); fail("host_alloc(%zu) failed", bytes); return nullptr; }
    return p;
}
void gpu_codec_host_free(void* p) { if (p) cudaFreeHost(p); }
void gpu_codec_free(void* p) { if (p && !pool_put(p)) free(p); }

// ------------------------------------------------------------------ device-resident

bool gpu_codec_encode_bgr_dev(int ctx, const uint8_t* d_rows, int width, int height, long long row_stride,
                              int window_y0, int window_y1, int mb_row0, int mb_row1, uint8_t* d_payload,
                              size_t payload_cap, uint64_t* d_payload_len, uint32_t* d_chunk_off, int8_t* d_qcoef,
                              void* stream) {
    Ctx* c = get_ctx(ctx);
    if (!c) return false;
    cudaStream_t st = (cudaStream_t) stream;
    return encode_dev_impl(c, d_rows, width, height, row_stride, window_y0, window_y1, mb_row0, mb_row1, d_payload,
                           payload_cap, d_payload_len, d_chunk_off, d_qcoef, st);
}

bool gpu_codec_encode_batch_bgr_dev(int ctx, const uint8_t* d_frames, int n_frames, long long frame_stride, int width,
                                    int height, long long row_stride, uint8_t* d_payload, size_t payload_cap,
                                    uint64_t* d_frame_off, void* stream) {
    Ctx* c = get_ctx(ctx);
    if (!c) return false;
    if (n_frames < 1 || frame_stride < row_stride * (long long) height) return fail("encode_batch: bad frame count or frame stride");
    if (!d_frame_off) return fail("encode_batch: null device pointer");
    const int mbh = (height + 15) / 16;
    return encode_dev_impl(c, d_frames, width, height, row_stride, 0, height, 0, mbh, d_payload, payload_cap, nullptr, nullptr,
                           nullptr, (cudaStream_t) stream, nullptr, 0, n_frames, frame_stride,
                           reinterpret_cast<unsigned long long*>(d_frame_off));
}

bool gpu_codec_decode_dev(int ctx, const uint8_t* d_payload, size_t len, int width, int height, int mb_row0,
                          int mb_row1, uint8_t* d_dst, long long row_stride, int rgb_topdown,
                          const uint32_t* d_chunk_off, void* stream) {
    Ctx* c = get_ctx(ctx);
    if (!c) return false;
    cudaStream_t st = (cudaStream_t) stream;
    return decode_dev_impl(c, d_payload, len, width, height, mb_row0, mb_row1, d_dst, row_stride, rgb_topdown,
                           d_chunk_off, st);
}

bool gpu_codec_sync_check(int ctx, void* stream) {
    Ctx* c = get_ctx(ctx);
    if (!c) return false;
    return check_err(c, (cudaStream_t) stream, "sync_check");
}

bool gpu_codec_synth_bgr_dev(int ctx, uint8_t* d_rows, int width, int height, long long row_stride, int kind,
                             uint32_t seed, int frame, void* stream) {
    Ctx* c = get_ctx(ctx);
    if (!c) return false;
    if (width <= 0 || height <= 0 || row_stride < 3ll * width || kind < 0 || kind > 2) return fail("synth: bad arguments");
    CU(cudaSetDevice(c->dev));
    cudaStream_t st = (cudaStream_t) stream;
    synth_kernel<<<148 * 16, 256, 0, st>>>(d_rows, width, height, row_stride, kind, seed, frame);
    if (row_stride > 3ll * width) pad_zero_kernel<<<148, 256, 0, st>>>(d_rows, width, height, row_stride);
    CU(cudaGetLastError());
    return true;
}

void gpu_codec_selftest_tile_buffer(int bytes) { g_selftest_compact_cap = bytes > 0 ? bytes : 0; }

bool gpu_codec_selftest_colour_dev(int ctx, const uint8_t* d_in, size_t n, int inverse, uint8_t* d_out, void* stream) {
    Ctx* c = get_ctx(ctx);
    if (!c) return false;
    CU(cudaSetDevice(c->dev));
    cudaStream_t st = (cudaStream_t) stream;
    colour_kernel<<<148 * 8, 256, 0, st>>>(d_in, n, inverse, d_out);
    CU(cudaGetLastError());
    return true;
}

// ------------------------------------------------------------------ host-buffer paths

// Encode macroblock rows [R0,R1) of a host frame on context c, leaving the payload in c->out and
// its length in *c->h_len after the compute stream drains.  The stripe's pixel rows are copied in
// sub-stripes on the copy stream; each sub-stripe's kernels start as soon as its rows (and the
// halo rows above them) have landed, so the PCIe transfer hides the compute.
static bool encode_stripe_async(Ctx* c, const uint8_t* bmp_rows, int width, int height, int row_stride, int R0, int R1,
                                int* n_sub_out = nullptr) {
    const int mbw = (width + 15) / 16;
    const int y_lo = R0 > 0 ? 16 * R0 - 16 : 0, y_hi = 16 * R1 < height ? 16 * R1 : height;
    const size_t in_bytes = (size_t) (y_hi - y_lo) * row_stride;
    const size_t cap = (size_t) (R1 - R0) * mbw * 6 * 121 + 16;
    CU(cudaSetDevice(c->dev));
    if (!ensure(c->in, in_bytes) || !ensure(c->out, cap)) return false;
    // ~48 MB per sub-stripe for large frames; smaller frames are still cut in about eight (not below 4 MB a piece)
    // so that most of their encode time hides behind the upload too
    size_t sub_bytes = in_bytes / 8;
    if (sub_bytes > SUB_BYTES) sub_bytes = SUB_BYTES;
    if (sub_bytes < SUB_BYTES_MIN) sub_bytes = SUB_BYTES_MIN;
    int n_sub = (int) ((in_bytes + sub_bytes - 1) / sub_bytes);
    if (n_sub > MAX_SUB) n_sub = MAX_SUB;
    if (n_sub > R1 - R0) n_sub = R1 - R0;
    if (n_sub < 1) n_sub = 1;
    CU(cudaMemsetAsync(c->d_run, 0, 16, c->stream));
    const bool pageable = is_pageable(bmp_rows);
    int y_done = y_lo;                                   // rows [y_lo, y_done) are on their way
    for (int s = 0; s < n_sub; s++) {
        const int r0 = R0 + (int) ((long long) (R1 - R0) * s / n_sub), r1 = R0 + (int) ((long long) (R1 - R0) * (s + 1) / n_sub);
        const int y_to = 16 * r1 < height ? 16 * r1 : height;
        if (y_to > y_done) {
            // rows y_done..y_to-1 are the bottom-up byte range [(height-y_to)*stride, (height-y_done)*stride)
            const uint8_t* hsrc = bmp_rows + (size_t) (height - y_to) * row_stride;
            uint8_t* ddst = c->in.p + (size_t) (y_hi - y_to) * row_stride;
            if (!upload(c, ddst, hsrc, (size_t) (y_to - y_done) * row_stride, pageable)) return false;
            y_done = y_to;
        }
        CU(cudaEventRecord(c->ev[s], c->copy_stream));
        CU(cudaStreamWaitEvent(c->stream, c->ev[s], 0));
        if (!encode_dev_impl(c, c->in.p, width, height, row_stride, y_lo, y_hi, r0, r1, c->out.p, cap, nullptr, nullptr,
                             nullptr, c->stream, c->d_run, s))
            return false;
        CU(cudaMemcpyAsync(c->h_chain + s, c->d_run + ((s & 1) ^ 1), 8, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaEventRecord(c->ev_done[s], c->stream));
    }
    CU(cudaMemcpyAsync(c->h_len, c->d_run + (n_sub & 1), 8, cudaMemcpyDeviceToHost, c->stream));
    if (n_sub_out) *n_sub_out = n_sub;
    return true;
}

// Single-device frame: ship each sub-stripe's payload bytes home as soon as its kernels are done, on a third
// stream, while later sub-stripes are still uploading (PCIe is full duplex).  The result block is sized from
// the first sub-stripe's yield; if the guess turns out too small the payload (still complete on the device) is
// fetched again in one piece.
static bool encode_frame_streamed(Ctx* c, const uint8_t* bmp_rows, int width, int height, int row_stride,
                                  uint8_t** payload, size_t* payload_len) {
    const int mbh = (height + 15) / 16;
    int n_sub = 1;
    if (!encode_stripe_async(c, bmp_rows, width, height, row_stride, 0, mbh, &n_sub)) { drain(c); return false; }
    uint8_t* buf = nullptr;
    size_t cap = 0, prev = 0;
    bool overflow = false, ok = true;
    for (int s = 0; s < n_sub && ok; s++) {
        cudaError_t e = cudaEventSynchronize(c->ev_done[s]);
        if (e != cudaSuccess) { ok = fail("encode: waiting for sub-stripe %d failed: %s", s, cudaGetErrorString(e)); break; }
        const size_t cur = (size_t) c->h_chain[s];
        if (s == 0) {
            cap = cur * (size_t) n_sub + cur * (size_t) n_sub / 2 + (1u << 20);
            buf = static_cast<uint8_t*>(pool_get(cap));
            if (!buf) { ok = fail("encode: out of page-locked host memory (%zu bytes)", cap); break; }
        }
        if (cur > cap || cur < prev) overflow = true;
        if (!overflow && cur > prev) {
            e = cudaMemcpyAsync(buf + prev, c->out.p + prev, cur - prev, cudaMemcpyDeviceToHost, c->back_stream);
            if (e != cudaSuccess) ok = fail("encode: D2H failed: %s", cudaGetErrorString(e));
        }
        prev = cur;
    }
    if (!check_err(c, c->stream, "encode")) ok = false;      // also drains the compute stream on the failure paths
    const size_t total = ok ? (size_t) *c->h_len : 0;
    // the block may only change hands once nothing is being copied into it any more
    if (cudaStreamSynchronize(c->back_stream) != cudaSuccess && ok) ok = fail("encode: D2H failed");
    if (ok && overflow) {
        pool_put(buf);
        buf = static_cast<uint8_t*>(pool_get(total));
        if (!buf) return fail("encode: out of page-locked host memory (%zu bytes)", total);
        cudaError_t e = cudaMemcpyAsync(buf, c->out.p, total, cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) ok = fail("encode: D2H failed: %s", cudaGetErrorString(e));
    }
    if (!ok) { drain(c); if (buf) pool_put(buf); return false; }
    *payload = buf;
    *payload_len = total;
    return true;
}

bool gpu_codec_encode_bgr(const uint8_t* bmp_rows, int width, int height, int row_stride, uint8_t** payload,
                          size_t* payload_len) {
    if (!payload || !payload_len) return fail("encode: null output argument");
    *payload = nullptr;
    *payload_len = 0;
    if (g_ctx.empty()) return fail("encode: not initialised (call gpu_codec_init)");
    if (!bmp_rows || width <= 0 || height <= 0 || row_stride < 3ll * width) return fail("encode: bad arguments");
    const int mbh = (height + 15) / 16;
    const int G = (int) g_ctx.size() < mbh ? (int) g_ctx.size() : mbh;
    std::vector<std::unique_lock<std::mutex>> locks;
    for (int g = 0; g < G; g++) locks.emplace_back(g_ctx[g]->mu);
    if (G == 1) return encode_frame_streamed(g_ctx[0], bmp_rows, width, height, row_stride, payload, payload_len);

    // phase 1: every GPU gets its stripe (+ the macroblock row above it) and encodes it
    for (int g = 0; g < G; g++) {
        int r0, r1;
        stripe_rows(mbh, g, G, r0, r1);
        if (!encode_stripe_async(g_ctx[g], bmp_rows, width, height, row_stride, r0, r1)) {
            for (int k = 0; k <= g; k++) drain(g_ctx[k]);
            return false;
        }
    }
    // phase 2: exclusive prefix sum of the per-stripe byte lengths, stitch (byte seams)
    size_t total = 0;
    std::vector<size_t> off(G), len(G);
    for (int g = 0; g < G; g++) {
        if (!check_err(g_ctx[g], g_ctx[g]->stream, "encode")) {
            for (int k = 0; k < G; k++) drain(g_ctx[k]);
            return false;
        }
        off[g] = total;
        len[g] = (size_t) *g_ctx[g]->h_len;
        total += len[g];
    }
    uint8_t* buf = static_cast<uint8_t*>(pool_get(total));
    if (!buf) return fail("encode: out of page-locked host memory (%zu bytes)", total);
    for (int g = 0; g < G; g++) {
        Ctx* c = g_ctx[g];
        cudaSetDevice(c->dev);
        cudaError_t e = cudaMemcpyAsync(buf + off[g], c->out.p, len[g], cudaMemcpyDeviceToHost, c->stream);
        if (e != cudaSuccess) {
            for (int k = 0; k < G; k++) drain(g_ctx[k]);
            pool_put(buf);
            return fail("encode: D2H failed: %s", cudaGetErrorString(e));
        }
    }
    for (int g = 0; g < G; g++) {
        cudaSetDevice(g_ctx[g]->dev);
        cudaError_t e = cudaStreamSynchronize(g_ctx[g]->stream);
        if (e != cudaSuccess) {
            for (int k = 0; k < G; k++) drain(g_ctx[k]);
            pool_put(buf);
            return fail("encode: sync failed: %s", cudaGetErrorString(e));
        }
    }
    *payload = buf;
    *payload_len = total;
    return true;
}

// One device's share of a host batch: frames i = first, first+step, ...  They are copied in groups of
// ~SUB_BYTES on the copy stream; each group is encoded by ONE set of launches (batch kernel) as soon as it
// has landed; payloads accumulate in c->out with their offsets in d_off.
static bool encode_batch_on_ctx(Ctx* c, const uint8_t* const* frames, int n_frames, int first, int step, int width, int height,
                                int row_stride, uint8_t** payloads, size_t* payload_lens) {
    const int mbw = (width + 15) / 16, mbh = (height + 15) / 16;
    const int n_mine = first < n_frames ? (n_frames - first + step - 1) / step : 0;
    if (n_mine == 0) return true;
    const size_t frame_bytes = (size_t) height * row_stride;
    const size_t fs = (frame_bytes + 255) / 256 * 256;                 // device frame stride
    const size_t cap = (size_t) n_mine * ((size_t) mbw * mbh * 6 * 121) + 16;
    CU(cudaSetDevice(c->dev));
    if (!ensure(c->in, fs * n_mine) || !ensure(c->out, cap) || !ensure(c->frame_off, (size_t) (n_mine + 1) * 8)) return false;
    unsigned long long* d_off = reinterpret_cast<unsigned long long*>(c->frame_off.p);   // n_mine + 1 offsets (+ pinned host copy)
    unsigned long long* h_off = static_cast<unsigned long long*>(pool_get((size_t) (n_mine + 1) * 8));
    bool ok = h_off != nullptr;
    if (!ok) fail("encode_batch: out of page-locked host memory");
    int per_group = (int) (SUB_BYTES / fs);
    if (per_group < 1) per_group = 1;
    // Groups are queued one ahead: while group g uploads, the host waits for group g-1's kernels (they finish right
    // behind its upload), reads that group's frame offsets, takes a result block per frame and queues the payload
    // copies on the third stream (PCIe is full duplex).
    auto ship = [&](int f0, int f1, cudaEvent_t done) -> bool {
        CU(cudaEventSynchronize(done));
        for (int f = f0; f < f1; f++) {
            const size_t n = (size_t) (h_off[f + 1] - h_off[f]);
            const int i = first + f * step;
            payloads[i] = static_cast<uint8_t*>(pool_get(n));
            if (!payloads[i]) return fail("encode_batch: out of page-locked host memory");
            payload_lens[i] = n;
            CU(cudaMemcpyAsync(payloads[i], c->out.p + h_off[f], n, cudaMemcpyDeviceToHost, c->back_stream));
        }
        return true;
    };
    auto body = [&]() -> bool {
        CU(cudaMemsetAsync(c->d_run, 0, 16, c->stream));
        int gi = 0, prev_f0 = -1, prev_f1 = -1;
        cudaEvent_t prev_done = nullptr;
        for (int f0 = 0; f0 < n_mine; f0 += per_group, gi++) {
            const int f1 = f0 + per_group < n_mine ? f0 + per_group : n_mine;
            for (int f = f0; f < f1; f++) {
                const uint8_t* src = frames[first + (size_t) f * step];
                if (!upload(c, c->in.p + (size_t) f * fs, src, frame_bytes, is_pageable(src))) return false;
            }
            cudaEvent_t ev = c->ev[gi % MAX_SUB], done = c->ev_done[gi % MAX_SUB];
            CU(cudaEventRecord(ev, c->copy_stream));
            CU(cudaStreamWaitEvent(c->stream, ev, 0));
            if (!encode_dev_impl(c, c->in.p + (size_t) f0 * fs, width, height, row_stride, 0, height, 0, mbh, c->out.p, cap, nullptr,
                                 nullptr, nullptr, c->stream, c->d_run, gi, f1 - f0, (long long) fs, d_off + f0))
                return false;
            CU(cudaMemcpyAsync(h_off + f0, d_off + f0, (size_t) (f1 - f0 + 1) * 8, cudaMemcpyDeviceToHost, c->stream));
            CU(cudaEventRecord(done, c->stream));
            if (prev_done && !ship(prev_f0, prev_f1, prev_done)) return false;
            prev_f0 = f0; prev_f1 = f1; prev_done = done;
        }
        if (prev_done && !ship(prev_f0, prev_f1, prev_done)) return false;
        if (!check_err(c, c->stream, "encode_batch")) return false;
        CU(cudaStreamSynchronize(c->back_stream));
        return true;
    };
    ok = ok && body();
    if (!ok) { cudaStreamSynchronize(c->stream); cudaStreamSynchronize(c->back_stream); }   // nothing may still be copying into blocks the caller is about to release
    if (h_off) pool_put(h_off);
    return ok;
}

bool gpu_codec_encode_batch_bgr(const uint8_t* const* frames, int n_frames, int width, int height, int row_stride,
                                uint8_t** payloads, size_t* payload_lens) {
    if (g_ctx.empty()) return fail("encode_batch: not initialised (call gpu_codec_init)");
    if (!frames || !payloads || !payload_lens || n_frames < 0 || width <= 0 || height <= 0 || row_stride < 3ll * width)
        return fail("encode_batch: bad arguments");
    for (int i = 0; i < n_frames; i++) { payloads[i] = nullptr; payload_lens[i] = 0; }
    const int G = (int) g_ctx.size();
    std::vector<std::string> errs(G);
    auto run = [&](int g) {
        Ctx* c = g_ctx[g];
        std::lock_guard<std::mutex> lk(c->mu);
        if (!encode_batch_on_ctx(c, frames, n_frames, g, G, width, height, row_stride, payloads, payload_lens)) errs[g] = g_last_error;
    };
    if (G == 1) {
        run(0);
    } else {
        std::vector<std::thread> workers;
        for (int g = 0; g < G; g++) workers.emplace_back(run, g);
        for (auto& t : workers) t.join();
    }
    for (int g = 0; g < G; g++)
        if (!errs[g].empty()) {
            for (int i = 0; i < n_frames; i++) { gpu_codec_free(payloads[i]); payloads[i] = nullptr; payload_lens[i] = 0; }
            return fail("%s", errs[g].c_str());
        }
    return true;
}

static bool decode_host(const uint8_t* payload, size_t len, int width, int height, uint8_t* dst, long long stride, int rgb) {
    if (g_ctx.empty()) return fail("decode: not initialised (call gpu_codec_init)");
    if (!payload || !dst || width <= 0 || height <= 0) return fail("decode: bad arguments");
    const int mbh = (height + 15) / 16;
    const int G = (int) g_ctx.size() < mbh ? (int) g_ctx.size() : mbh;
    const long long pitch = rgb ? 3ll * width : stride;
    std::vector<std::unique_lock<std::mutex>> locks;
    for (int g = 0; g < G; g++) locks.emplace_back(g_ctx[g]->mu);
    const bool dst_pageable = is_pageable(dst) && pitch == 3ll * width, src_pageable = is_pageable(payload) && len >= (4u << 20);
    std::vector<std::vector<PendingOut>> pend(G);
    auto enqueue = [&](int g) -> bool {
        Ctx* c = g_ctx[g];
        int R0, R1;
        stripe_rows(mbh, g, G, R0, R1);
        const int y_lo = 16 * R0, y_hi = 16 * R1 < height ? 16 * R1 : height;
        const size_t out_bytes = (size_t) (y_hi - y_lo) * pitch;
        CU(cudaSetDevice(c->dev));
        if (!ensure(c->in, len + 16) || !ensure(c->out, out_bytes)) return false;
        if (src_pageable) {
            if (!upload(c, c->in.p, payload, len, true)) return false;
            CU(cudaEventRecord(c->ev_done[0], c->copy_stream));
            CU(cudaStreamWaitEvent(c->stream, c->ev_done[0], 0));
        } else {
            CU(cudaMemcpyAsync(c->in.p, payload, len, cudaMemcpyHostToDevice, c->stream));
        }
        // the stripe's rows are contiguous in both sinks; hand the kernel the address the
        // full frame's first byte would have
        uint8_t* vbase = rgb ? c->out.p - (long long) y_lo * pitch : c->out.p - (long long) (height - y_hi) * pitch;
        int n_sub = (int) ((out_bytes + SUB_BYTES - 1) / SUB_BYTES);
        if (n_sub > MAX_SUB) n_sub = MAX_SUB;
        if (n_sub > R1 - R0) n_sub = R1 - R0;
        if (n_sub < 1) n_sub = 1;
        for (int s = 0; s < n_sub; s++) {
            // parse once (first call), then reconstruct sub-stripes and ship each while the next one runs
            const int r0 = R0 + (int) ((long long) (R1 - R0) * s / n_sub), r1 = R0 + (int) ((long long) (R1 - R0) * (s + 1) / n_sub);
            if (!decode_dev_impl(c, c->in.p, len, width, height, r0, r1, vbase, pitch, rgb, nullptr, c->stream, s > 0)) return false;
            CU(cudaEventRecord(c->ev[s], c->stream));
            CU(cudaStreamWaitEvent(c->copy_stream, c->ev[s], 0));
            const int ya = 16 * r0, yb = 16 * r1 < height ? 16 * r1 : height;
            const size_t rel = rgb ? (size_t) (ya - y_lo) * pitch : (size_t) (y_hi - yb) * pitch;      // offset inside the stripe buffer
            uint8_t* hdst = rgb ? dst + (size_t) ya * pitch : dst + (size_t) (height - yb) * pitch;
            if (dst_pageable) {
                if (!download(c, hdst, c->out.p + rel, (size_t) (yb - ya) * pitch, pend[g])) return false;
            } else if (pitch == 3ll * width) {
                CU(cudaMemcpyAsync(hdst, c->out.p + rel, (size_t) (yb - ya) * pitch, cudaMemcpyDeviceToHost, c->copy_stream));
            } else {   // leave the caller's row padding bytes alone
                CU(cudaMemcpy2DAsync(hdst, (size_t) pitch, c->out.p + rel, (size_t) pitch, (size_t) 3 * width, (size_t) (yb - ya),
                                     cudaMemcpyDeviceToHost, c->copy_stream));
            }
        }
        return true;
    };
    bool ok = true;
    for (int g = 0; g < G && ok; g++) ok = enqueue(g);
    // success or not, nothing may still be writing into the caller's buffer when this returns
    for (int g = 0; g < G; g++) {
        cudaSetDevice(g_ctx[g]->dev);
        while (!pend[g].empty())
            if (!drain_one(g_ctx[g], pend[g])) { ok = false; pend[g].clear(); }
        if (!check_err(g_ctx[g], g_ctx[g]->stream, "decode")) ok = false;
        cudaSetDevice(g_ctx[g]->dev);
        if (cudaStreamSynchronize(g_ctx[g]->copy_stream) != cudaSuccess) ok = fail("decode: D2H failed");
    }
    return ok;
}

bool gpu_codec_decode_to_bgr(const uint8_t* payload, size_t len, int width, int height, uint8_t* bmp_rows, int row_stride) {
    if (row_stride < 3ll * width) return fail("decode: row_stride %d < 3*width", row_stride);
    return decode_host(payload, len, width, height, bmp_rows, row_stride, 0);
}

bool gpu_codec_decode_to_rgb24(const uint8_t* payload, size_t len, int width, int height, uint8_t* rgb_topdown) {
    return decode_host(payload, len, width, height, rgb_topdown, 3ll * width, 1);
}

}  // extern "C"

// ------------------------------------------------------------------ encode without allocating
//
// Some callers may not allocate device memory at the point where they encode: a frame
// loop that has handed the whole device to its own pool, a latency-critical path that
// cannot take the device-wide synchronisation a cudaMalloc/cudaFree implies, or code that
// runs while some other stream is being captured into a graph (allocation calls are not
// allowed then).  The regular device entry point grows the context's scratch with
// cudaMalloc on demand, so it is not for them.  This variant keeps the encode kernel's
// scratch — the tile ticket and one look-back descriptor per 32-macroblock tile — in a
// fixed array that is part of the library image.  Every GPU that loads the library
// carries these 512 KB + 256 B until the process exits.
namespace {

constexpr int NOALLOC_MAX_TILES = 2097152 / awi::ENC_TILE_MB;   // 65536 tiles = 2,097,152 macroblocks, e.g. up to 23168 x 23168 pixels
static_assert(awi::ENC_TILE_MB == 32, "include/gpu_codec.h, DESIGN.md and INTEGRATION.md quote 65536 tiles / 512 KB + 256 B for 32-macroblock tiles");
constexpr size_t NOALLOC_HEADER = 256;            // [0,8) spare, [8,12) tile ticket
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_HEADER + (size_t) NOALLOC_MAX_TILES * 8];

// In host code `g_noalloc_scratch` is only a handle for the runtime, not an address the
// GPU understands; the real per-device address comes from cudaGetSymbolAddress, asked
// once per device with that device current, and is what the kernel is given.
constexpr int NOALLOC_MAX_DEVICES = 64;
unsigned char* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = { nullptr };
std::mutex g_noalloc_mu;

bool noalloc_scratch_for_current_device(unsigned char** out) {
    int dev = -1;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return fail("encode_noalloc: device index %d not supported", dev);
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) return fail("encode_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
        g_noalloc_ptr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_noalloc_ptr[dev];
    return true;
}

}  // namespace

extern "C" bool gpu_codec_encode_bgr_dev_noalloc(const uint8_t* d_rows, int width, int height, long long row_stride,
                                                 uint8_t* d_payload, size_t payload_cap, uint64_t* d_payload_len,
                                                 uint32_t* d_status, void* stream) {
    if (width <= 0 || height <= 0) return fail("encode_noalloc: bad geometry %dx%d", width, height);
    if (row_stride < 3ll * width) return fail("encode_noalloc: row_stride %lld < 3*width", row_stride);
    if (!d_rows || !d_payload || !d_payload_len || !d_status) return fail("encode_noalloc: null device pointer");
    const int mbw = (width + 15) / 16, mbh = (height + 15) / 16;
    const long long n_mb = (long long) mbw * mbh;
    const long long n_tiles = (n_mb + awi::ENC_TILE_MB - 1) / awi::ENC_TILE_MB;
    if (n_tiles > NOALLOC_MAX_TILES)
        return fail("encode_noalloc: %dx%d needs %lld tiles of %d macroblocks, the built-in scratch holds %d "
                    "(at most %lld macroblocks per call); use gpu_codec_encode_bgr_dev for larger frames",
                    width, height, n_tiles, awi::ENC_TILE_MB, NOALLOC_MAX_TILES, (long long) NOALLOC_MAX_TILES * awi::ENC_TILE_MB);
    unsigned char* scratch = nullptr;
    if (!noalloc_scratch_for_current_device(&scratch)) return false;
    cudaStream_t st = (cudaStream_t) stream;
    // the array still holds the previous call's ticket and descriptors: both must start at zero
    CU(cudaMemsetAsync(scratch, 0, NOALLOC_HEADER + (size_t) n_tiles * 8, st));
    CU(cudaMemsetAsync(d_status, 0, 4, st));

    awi::EncParams P{};
    P.rows = d_rows;
    P.stride = row_stride;
    P.width = width;
    P.height = height;
    P.align = align_of(d_rows, row_stride);
    P.mbw = mbw;
    P.mb0 = 0;
    P.mb1 = (int) n_mb;
    P.out = d_payload;
    P.out_cap = payload_cap;
    P.out_len = reinterpret_cast<unsigned long long*>(d_payload_len);
    P.ticket = reinterpret_cast<unsigned int*>(scratch + 8);
    P.desc = reinterpret_cast<unsigned long long*>(scratch + NOALLOC_HEADER);
    P.err = d_status;
    P.chunk_off = nullptr;
    P.qcoef = nullptr;
    P.compact_cap = g_selftest_compact_cap;
    CU(awi::launch_encode(P, (int) n_tiles, st));
    return true;
}
