#!/usr/bin/env python3
"""bench.py — Mpixel/s of the AWI hot path on B200, one JSON line on stdout.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                  [--workload c3|c2|c4|c5] [--kind photo|noise]

Workloads restate BASELINE.json `configs`:
  c3 (default) configs[2]  encode one synthetic 16384x16384 frame.  N > 1: THE SAME frame cut into N
               macroblock-row stripes, a stripe per GPU (strong scaling, no collective: chunks are byte aligned,
               rle.c:107, and the six predictors, awi.c:91-92, are rebuilt from a halo row); the stripes are
               byte-verified against the 1-GPU payload (sha256 per segment).
  c2           configs[1]  encode one 3840x2160 frame (12 frames in rotation so the input is cold in L2).
  c4           configs[3]  encode a batch of 1920x1080 frames, 128 per GPU (1024 at 8 GPUs), whole frames per GPU.
  c5           configs[4]  stand-alone decode of the .awi payload of an 8192x8192 frame (boundary discovery, DC
               scans, reconstruction).  N > 1: replicated parse, sharded reconstruct (rows of stripe r on GPU r),
               pixels byte-verified against the 1-GPU decode.

A "step" is one pass of the hot path over one batch of input resident in HBM.
`value`     device-resident throughput: CUDA events on the launching stream, max over ranks.
`e2e`       the same metric through the reference-facing C ABI with pinned HOST buffers, H2D and D2H inside the
            timed region: gpu_codec_encode_bgr / gpu_codec_encode_batch_bgr / gpu_codec_decode_to_bgr.  At N > 1
            for the striped workloads (c3, c5) it is the in-process multi-GPU path a drop-in host would use:
            rank 0 alone calls gpu_codec_init_devices([0..N-1]) and then the same entry point on ONE host frame
            (stripes + host stitch from the prefix sum of the N lengths), the other ranks idle; the result's
            sha256 must equal the 1-GPU result's.
`roofline`  algorithmic bytes (3WH + payload) / step time against the measured HBM copy peak (x N GPUs).
`cpu_baseline` the reference's own C code (oracle/_ref, one thread: it has none) on a bounded sample.
`--impl reference` that code on all host cores (one band per thread), same config dict.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import statistics
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # op, width, height, frames per GPU, how N GPUs share the work, input copies rotated through, description
    "c3": dict(op="encode", w=16384, h=16384, frames=1, split="stripes", rotate=1,
               desc="configs[2]: synthetic 16384x16384 BGR frame encode, macroblock-row stripes over the GPUs"),
    "c2": dict(op="encode", w=3840, h=2160, frames=1, split="frames", rotate=12,
               desc="configs[1]: synthetic 3840x2160 BGR single-frame encode"),
    "c4": dict(op="encode", w=1920, h=1080, frames=128, split="frames", rotate=1,
               desc="configs[3]: batch of 1920x1080 frames, 128 per GPU (1024 at 8 GPUs), image-parallel"),
    "c5": dict(op="decode", w=8192, h=8192, frames=1, split="stripes", rotate=24,
               desc="configs[4]: round-trip decode of the .awi of a synthetic 8192x8192 frame, stripes over the GPUs"),
}
KIND = {"noise": 0, "photo": 1, "flat": 2}
CFG_NO = {"c2": 2, "c3": 3, "c4": 4, "c5": 5}
L2_BYTES = 126 << 20
DTYPE = "f32+f64 (bit-exact u8 in/out)"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def make_config(args, world: int) -> dict:
    """The `config` object — built by one function so that both arms print the same keys and values."""
    W = WORKLOADS[args.workload]
    w, h, frames = W["w"], W["h"], W["frames"]
    frame_bytes = 3 * w * h
    if W["split"] == "stripes" and world > 1:
        par = (f"one frame in {world} macroblock-row stripes, a stripe per GPU; no collective "
               f"(host stitch from the prefix sum of {world} byte lengths)" if W["op"] == "encode" else
               f"replicated parse, reconstruction of one frame sharded in {world} macroblock-row stripes; no collective")
    else:
        par = f"{world} x independent frames, no collective" if world > 1 else "single GPU"
    if W["op"] == "encode":
        per_step = frame_bytes * frames
        l2 = ("inputs larger than L2 (%.0f MB read per step vs 126 MB), no flush needed" % (per_step / 1e6)
              if per_step > 2 * L2_BYTES else
              "a frame fits L2: %d distinct frames in rotation (%.0f MB), each step reads one that has been evicted"
              % (W["rotate"], W["rotate"] * per_step / 1e6))
    else:
        l2 = ("%.0f MB of pixels written per step (larger than L2); the payload rotates through %d copies so it is "
              "read from HBM too" % (frame_bytes / 1e6, W["rotate"]))
    return {"workload": W["desc"], "op": W["op"], "kind": args.kind, "width": w, "height": h, "frames_per_gpu": frames,
            "l2": l2, "parallelism": par}


class ClockSampler:
    """SM clock + throttle reasons sampled through NVML while the timed region runs."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake", 0x2: "applications_clocks_setting", 0x100: "display_clock_setting"}

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self._h = None
        self._t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self._nv.nvmlDeviceGetClockInfo(self._h, self._nv.NVML_CLOCK_SM))
                r = self._nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.004)

    def __enter__(self):
        if self._h is not None:
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._h is not None:
            self._t.join()

    def summary(self):
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def _seed(workload: str) -> int:
    from image_compression_b200 import synth
    return synth.SEED_BASE + CFG_NO[workload]


def sha(b) -> str:
    return hashlib.sha256(memoryview(np.ascontiguousarray(b))).hexdigest()


# ----------------------------------------------------------------------------- reference arm / CPU baseline
# (the only places bench.py touches oracle/: the reference's own C code as the timed CPU baseline)

def _reference_lib():
    import oracle
    r = oracle.ref()
    if r is not None:
        return r, "reference"
    return oracle.port(), "port"


def _band_rows(width: int, band_rows: int, kind: str, seed: int) -> np.ndarray:
    """The top `band_rows` rows of the workload frame (same generator, same seed: the first rows of the
    frame are the first rows of the band)."""
    from image_compression_b200 import synth
    return synth.frame_bgr_bottomup(kind, width, band_rows, seed, 0)


class CpuSample:
    """A bounded sample of the workload for the CPU arm: a band of the frame, wrapped the way the reference reads
    it (a BMP file image for compress_to_awi, an .awi file image for decompress_awi)."""

    def __init__(self, op: str, width: int, band: int, kind: str, seed: int):
        import oracle
        self.lib, self.kind_name = _reference_lib()
        self.op, self.width, self.band = op, width, band
        rows = _band_rows(width, band, kind, seed)
        self.bmp = np.frombuffer(oracle.bmp_wrap(rows, width, band), np.uint8)
        if op == "decode":          # the stream to decode is made by the same CPU code, outside the timed region
            if self.kind_name == "reference":
                self.awi = self.lib.encode_bmp(self.bmp)
            else:
                self.payload = self.lib.encode_bgr(self.bmp[54:], width, band, 3 * width)

    def run_once(self) -> float:
        t = time.perf_counter()
        if self.op == "encode":
            if self.kind_name == "reference":
                self.lib.encode_bmp(self.bmp)
            else:
                self.lib.encode_bgr(self.bmp[54:], self.width, self.band, 3 * self.width)
        else:
            if self.kind_name == "reference":
                self.lib.decode_to_bmp(self.awi)
            else:
                self.lib.decode(self.payload, self.width, self.band, 3 * self.width)
        return time.perf_counter() - t

    def fn_name(self) -> str:
        return "compress_to_awi" if self.op == "encode" else "decompress_awi"


def cpu_baseline(args) -> dict:
    W = WORKLOADS[args.workload]
    w, h = W["w"], W["h"]
    band = min(h, 2048 if w >= 8192 else h)
    s = CpuSample(W["op"], w, band, args.kind, _seed(args.workload))
    dt = min(s.run_once() for _ in range(2))
    return {"value": round(w * band / dt / 1e6, 3), "unit": "Mpixel/s", "cores": 1, "kind": s.kind_name,
            "sample": f"top {band} rows of the {w}x{h} {args.kind} frame ({w * band / 1e6:.1f} Mpx), "
                      f"{s.fn_name()} from memory, gcc -O3 -ffp-contract=off, best of 2",
            "seconds": round(dt, 3)}


def run_reference_arm(args) -> dict:
    W = WORKLOADS[args.workload]
    w, h = W["w"], W["h"]
    cores = os.cpu_count() or 1
    band = 256 if w >= 8192 else min(h, 1088)
    s = CpuSample(W["op"], w, band, args.kind, _seed(args.workload))
    px_per_step = w * band * cores

    def step():
        ts = [threading.Thread(target=s.run_once) for _ in range(cores)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    val = round(px_per_step * args.steps / dt / 1e6, 3)
    sample = (f"{cores} threads x a {w}x{band} band of the {w}x{h} {args.kind} frame per step "
              f"({px_per_step / 1e6:.1f} Mpx/step); the reference's {s.fn_name()}, memory-backed")
    scaling = "strong" if (W["split"] == "stripes" and args.gpus > 1) else "weak"
    return {"impl": "reference", "metric": f"{W['op']}_mpixel_per_s", "value": val, "unit": "Mpixel/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3),
            "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": DTYPE,
            "data": "synthetic", "config": make_config(args, args.gpus),
            "cpu_baseline": {"value": val, "unit": "Mpixel/s", "cores": cores, "kind": s.kind_name, "sample": sample},
            "e2e": {"value": val, "unit": "Mpixel/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


# ----------------------------------------------------------------------------- our arm: helpers

def enc_launches(n_mb: int, frames: int = 1) -> int:
    """Kernels one device-resident encode call launches: awi_encode_kernel + the fused scan/gather pass up to
    65536 tiles of 32 macroblocks, otherwise + enc_scan_tiles + enc_gather_tiles (awi_encode.cu launch_encode)."""
    tiles = (n_mb + 31) // 32 * frames
    return 2 if tiles <= 65536 else 3


def dec_launches(payload_bytes: int, n_mb: int) -> int:
    """Kernels one stand-alone decode launches (awi_decode.cu launch_decode)."""
    n, levels = max(1, (payload_bytes + 2047) // 2048), 1
    while n > 1:
        n = (n + 31) // 32
        levels += 1
    front = 1 + (levels - 1) + 1 + (1 if levels >= 2 else 0) + 1      # segments, compose x (levels-1), root, expand, offsets
    dc = 2 if (n_mb + 255) // 256 <= 1024 else 3
    return front + dc + 1


class Env:
    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device — the codec has no CPU path (use --impl reference for the CPU arm)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.cpu_group = None
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
            self.cpu_group = dist.new_group(backend="gloo")      # host-side waits that must not occupy any GPU

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def host_barrier(self):
        if self.world > 1:
            self.dist.barrier(group=self.cpu_group)

    def max_ms(self, ms: float) -> float:
        t = self.torch.tensor([ms], dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def gather_obj(self, obj):
        if self.world == 1:
            return [obj]
        out = [None] * self.world
        self.dist.all_gather_object(out, obj, group=self.cpu_group)
        return out

    def timed(self, fn, steps: int, warmup: int, sample_clocks: bool = False):
        """warm-up, barrier + synchronize, K steps between two events on the current stream, barrier + synchronize;
        returns (ms per step on this rank, max over ranks of the total ms, clock summary)."""
        torch = self.torch
        for _ in range(warmup):
            fn()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        clk = ClockSampler(self.local) if sample_clocks else None
        if clk:
            clk.__enter__()
        ev0.record()
        for _ in range(steps):
            fn()
        ev1.record()
        self.barrier()
        if clk:
            clk.__exit__()
        ms = ev0.elapsed_time(ev1)
        return ms / steps, self.max_ms(ms), (clk.summary() if clk else None)

    def finish(self):
        if self.world > 1:
            self.dist.barrier()
            self.dist.destroy_process_group()


def pinned(lib, n: int):
    import ctypes as C
    p = lib.gpu_codec_host_alloc(n)
    if not p:
        raise SystemExit("bench.py: pinned allocation failed")
    return p, np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n,))


def traffic_from_profile(key: str):
    tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        d = json.load(open(tp))
        return d.get(key), d.get(key + "_source")
    except Exception:
        return None, None


# ----------------------------------------------------------------------------- our arm: encode workloads

def run_encode(args, E: Env) -> dict | None:
    import ctypes as C
    from image_compression_b200 import codec, stripes
    torch = E.torch
    W = WORKLOADS[args.workload]
    w, h, frames, rotate = W["w"], W["h"], W["frames"], W["rotate"]
    striped = W["split"] == "stripes" and E.world > 1
    stride = 3 * w
    frame_bytes = stride * h
    mbw, mbh = (w + 15) // 16, (h + 15) // 16
    seed, kind = _seed(args.workload), KIND[args.kind]
    codec.init([E.local])
    lib = codec.load()
    warm = max(args.warmup, 3)

    # resident inputs.  Striped: every rank holds frame 0.  Otherwise `frames` distinct frames per rank (frame
    # index unique across ranks), `rotate` such sets when one set would sit in L2.
    n_res = frames * rotate
    rows = torch.empty((n_res, frame_bytes), dtype=torch.uint8, device=E.dev)
    for f in range(n_res):
        fi = 0 if W["split"] == "stripes" else (E.rank * n_res + f)
        codec.synth_dev(0, rows[f], w, h, stride, kind, seed, fi)
    cap = codec.max_payload_bytes(w, h) + 16
    payload = torch.empty(cap * frames, dtype=torch.uint8, device=E.dev)
    plen = torch.zeros(1, dtype=torch.int64, device=E.dev)
    batch_off = torch.zeros(frames + 1, dtype=torch.int64, device=E.dev)
    codec.sync_check(0)

    # ---- the whole frame / batch on this GPU alone: the N = 1 headline, and at N > 1 the bytes to verify against
    it = [0]

    def step_whole():
        s = it[0] % rotate
        it[0] += 1
        if frames == 1:
            codec.encode_dev(0, rows[s], w, h, stride, payload, plen)
        else:      # image-parallel batch: one set of launches for all resident frames
            codec.encode_batch_dev(0, rows[s * frames:(s + 1) * frames], frames, frame_bytes, w, h, stride, payload, batch_off)

    ms_whole, ms_whole_max, clk = E.timed(step_whole, args.steps, warm, sample_clocks=True)
    codec.sync_check(0)
    it[0] = 0
    step_whole()
    codec.sync_check(0)
    payload_bytes = int(plen.item()) if frames == 1 else int(batch_off[-1].item())
    whole_host = payload[:payload_bytes].cpu().numpy()          # frame 0 (set 0): reference bytes for the checks below
    verify = {}
    extras = {}

    if striped:
        # ---- ONE frame, a stripe per rank (strong scaling).  Every rank encodes rows [r0, r1) from its own
        # window of the frame (the stripe + 16 halo rows) only.
        r0, r1 = stripes.stripe_rows(mbh, E.world)[E.rank]
        y0, y1 = stripes.stripe_window(h, r0, r1)
        b0, b1 = stripes.window_byte_range(h, stride, y0, y1)
        win = rows[0][b0:b1].clone()
        slen = torch.zeros(1, dtype=torch.int64, device=E.dev)

        def step_stripe():
            codec.encode_dev(0, win, w, h, stride, payload, slen, window=(y0, y1), mb_rows=(r0, r1))

        ms_mine, ms_max, clk = E.timed(step_stripe, args.steps, warm, sample_clocks=True)
        codec.sync_check(0)
        seg = payload[: int(slen.item())].cpu().numpy()
        parts = E.gather_obj((int(seg.size), sha(seg), round(ms_mine, 4)))
        lens = [p[0] for p in parts]
        offs = stripes.segment_offsets(lens)
        ok = [sha(whole_host[o:o + n]) == p[1] for o, n, p in zip(offs, lens, parts)]
        verify["stripe_segments_sha256_equal_1gpu_payload"] = bool(all(ok) and sum(lens) == payload_bytes)
        verify["stripe_bytes"] = lens
        extras["stripe_ms_per_rank"] = [p[2] for p in parts]
        extras["independent_frames_per_gpu"] = {      # round 1's headline, kept as context: no shared state, scales trivially
            "mpixel_per_s": round(w * h * E.world * args.steps / (ms_whole_max * 1e-3) / 1e6, 1), "scaling": "weak"}
        step_ms = ms_max / args.steps
        px_per_step = w * h
        alg_bytes = 3 * w * h + payload_bytes
        launches = args.steps * enc_launches((r1 - r0) * mbw)
        kernel_ms = ms_mine
    else:
        step_ms = ms_whole_max / args.steps
        px_per_step = w * h * frames * E.world
        alg_bytes = (3 * w * h * frames + payload_bytes) * E.world
        launches = args.steps * enc_launches(mbw * mbh, frames)
        kernel_ms = ms_whole
    value = px_per_step / (step_ms * 1e-3) / 1e6

    peak, peak_src = peaks()
    achieved = alg_bytes / (step_ms * 1e-3) / 1e9
    traffic, traffic_src = traffic_from_profile(f"{args.workload}_{args.kind}")
    n_tiles = (mbw * mbh + 31) // 32 * frames
    kernels = ("awi_encode_kernel + enc_scan_gather_tiles" if (n_tiles if not striped else n_tiles // E.world) <= 65536
               else "awi_encode_kernel + enc_scan_tiles + enc_gather_tiles")
    roofline = {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak * E.world, "unit": "GB/s",
                "frac": round(achieved / (peak * E.world), 4), "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": peak_src + (f" x {E.world} GPUs" if E.world > 1 else ""),
                "kernel": kernels + " (one encode step; the first is >95 % of it)", "kernel_ms": round(kernel_ms, 4),
                "algorithmic_bytes_per_launch": alg_bytes,
                "note": "integer/fp32 issue-bound under the bit-exactness rule, not DRAM-bound (DESIGN.md §4)"}

    # ---- end to end through the C ABI with pinned host buffers, H2D + D2H inside the timed region
    e2e_steps = max(2, min(args.steps, 8))
    E.barrier()
    if striped:
        # in-process multi-GPU: rank 0 drives all N GPUs through the drop-in call on ONE host frame
        e2e = None
        if E.rank == 0:
            hptr, host = pinned(lib, frame_bytes)
            torch.from_numpy(host).copy_(rows[0])
            torch.cuda.synchronize()
            codec.init(list(range(E.world)))
            try:
                def e2e_step():
                    addr, n = codec.encode_bgr_ptr(hptr, w, h, stride)
                    return addr, n
                for _ in range(2):
                    codec.free_payload(e2e_step()[0])
                t0 = time.perf_counter()
                for _ in range(e2e_steps - 1):
                    codec.free_payload(e2e_step()[0])
                addr, n = e2e_step()
                dt = time.perf_counter() - t0
                got = np.ctypeslib.as_array(C.cast(addr, C.POINTER(C.c_uint8)), shape=(n,))
                verify["e2e_stitched_payload_sha256_equal_1gpu_payload"] = bool(n == payload_bytes and sha(got) == sha(whole_host))
                verify["payload_sha256"] = sha(whole_host)
                codec.free_payload(addr)
                e2e = {"value": round(w * h * e2e_steps / dt / 1e6, 1), "unit": "Mpixel/s", "h2d_bytes_per_step": frame_bytes + 48 * w * (E.world - 1),
                       "d2h_bytes_per_step": n + 8 * E.world, "steps": e2e_steps,
                       "api": f"gpu_codec_init_devices([0..{E.world - 1}]) + gpu_codec_encode_bgr, one pinned host frame, "
                              f"{E.world} stripes stitched on the host (rank 0 drives all GPUs; wall clock)"}
            finally:
                codec.init([E.local])
                lib.gpu_codec_host_free(C.c_void_p(hptr))
        E.host_barrier()
    else:
        if frames > 1:        # image-parallel batch through gpu_codec_encode_batch_bgr: all frames from pinned memory
            hptr, hb = pinned(lib, frame_bytes * frames)
            torch.from_numpy(hb.reshape(frames, frame_bytes)).copy_(rows[:frames])
            torch.cuda.synchronize()
            ptrs = (C.c_void_p * frames)(*[hptr + f * frame_bytes for f in range(frames)])
            outs = (C.c_void_p * frames)()
            lens = (C.c_size_t * frames)()

            def e2e_step(check=False):
                if not lib.gpu_codec_encode_batch_bgr(ptrs, frames, w, h, stride, outs, lens):
                    raise SystemExit("bench.py: " + lib.gpu_codec_last_error().decode())
                tot = sum(lens[i] for i in range(frames))
                good = None
                if check:
                    cat = np.concatenate([np.ctypeslib.as_array(C.cast(outs[i], C.POINTER(C.c_uint8)), shape=(lens[i],))
                                          for i in range(frames)])
                    good = bool(cat.size == payload_bytes and sha(cat) == sha(whole_host))
                for i in range(frames):
                    lib.gpu_codec_free(C.c_void_p(outs[i]))
                return tot, good
        else:
            hptr, host = pinned(lib, frame_bytes)
            torch.from_numpy(host).copy_(rows[0])
            torch.cuda.synchronize()

            def e2e_step(check=False):
                addr, n = codec.encode_bgr_ptr(hptr, w, h, stride)
                good = None
                if check:
                    got = np.ctypeslib.as_array(C.cast(addr, C.POINTER(C.c_uint8)), shape=(n,))
                    good = bool(n == payload_bytes and sha(got) == sha(whole_host))
                codec.free_payload(addr)
                return n, good
        for _ in range(2):
            e2e_step()
        E.barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        ev0.record()
        for _ in range(e2e_steps):
            e2e_payload, _ = e2e_step()
        ev1.record()
        torch.cuda.synchronize()
        e2e_ms = E.max_ms(max(ev0.elapsed_time(ev1), (time.perf_counter() - t0) * 1e3))
        _, good = e2e_step(check=True)
        goods = E.gather_obj(good)
        verify["e2e_payload_sha256_equal_device_resident_payload"] = bool(all(goods))
        lib.gpu_codec_host_free(C.c_void_p(hptr))
        e2e = {"value": round(px_per_step * e2e_steps / (e2e_ms * 1e-3) / 1e6, 1), "unit": "Mpixel/s",
               "h2d_bytes_per_step": frame_bytes * frames, "d2h_bytes_per_step": e2e_payload + 8 * (frames + 1),
               "steps": e2e_steps,
               "api": ("gpu_codec_encode_bgr" if frames == 1 else "gpu_codec_encode_batch_bgr") + " (pinned host buffers"
                      + (", one process per GPU)" if E.world > 1 else ")")}

    if E.rank == 0 and E.world == 1 and not args.no_extras and args.workload == "c3":
        extras.update(run_extras(codec, torch, E.dev))

    out = None
    if E.rank == 0:
        cfg = make_config(args, E.world)
        out = {"metric": "encode_mpixel_per_s", "value": round(value, 1), "unit": "Mpixel/s", "n_gpus": E.world,
               "steps": args.steps, "warmup": warm, "ms_per_step": round(step_ms, 4),
               "higher_is_better": True, "scaling": "strong" if striped else "weak", "vs_baseline": None,
               "dtype": DTYPE, "data": "synthetic", "config": cfg, "payload_bytes": payload_bytes,
               "clocks": clk, "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "verify": verify}
        if E.world == 1:
            out["cpu_baseline"] = cpu_baseline(args)
        if extras:
            out["extras"] = extras
    return out


# ----------------------------------------------------------------------------- our arm: decode workload (c5)

def run_decode(args, E: Env) -> dict | None:
    import ctypes as C
    from image_compression_b200 import codec, stripes
    torch = E.torch
    W = WORKLOADS[args.workload]
    w, h, rotate = W["w"], W["h"], W["rotate"]
    striped = E.world > 1
    stride = 3 * w
    frame_bytes = stride * h
    mbw, mbh = (w + 15) // 16, (h + 15) // 16
    seed, kind = _seed(args.workload), KIND[args.kind]
    codec.init([E.local])
    lib = codec.load()
    warm = max(args.warmup, 3)

    # the stream to decode: this library's encoder output for frame 0 (bit-identical to the reference's file, tests/)
    rows = torch.empty(frame_bytes, dtype=torch.uint8, device=E.dev)
    codec.synth_dev(0, rows, w, h, stride, kind, seed, 0)
    tmp = torch.empty(codec.max_payload_bytes(w, h) + 16, dtype=torch.uint8, device=E.dev)
    plen = torch.zeros(1, dtype=torch.int64, device=E.dev)
    codec.encode_dev(0, rows, w, h, stride, tmp, plen)
    codec.sync_check(0)
    n = int(plen.item())
    pitch = (n + 16 + 255) // 256 * 256
    copies = torch.empty(rotate * pitch, dtype=torch.uint8, device=E.dev)
    for i in range(rotate):
        copies[i * pitch: i * pitch + n] = tmp[:n]
    del tmp
    dst = torch.zeros(frame_bytes, dtype=torch.uint8, device=E.dev)
    it = [0]

    def step_whole():
        s = it[0] % rotate
        it[0] += 1
        codec.decode_dev(0, copies[s * pitch: s * pitch + n], n, w, h, dst, stride)

    ms_whole, ms_whole_max, clk = E.timed(step_whole, args.steps, warm, sample_clocks=True)
    codec.sync_check(0)
    verify, extras = {}, {}
    whole_sha_rows = None
    if striped:
        r0, r1 = stripes.stripe_rows(mbh, E.world)[E.rank]
        y0, y1 = 16 * r0, min(h, 16 * r1)
        b0, b1 = stripes.window_byte_range(h, stride, y0, y1)
        mine_whole = sha(dst[b0:b1].cpu().numpy())          # my rows out of my own whole-frame decode
        dst.zero_()

        def step_stripe():
            s = it[0] % rotate
            it[0] += 1
            codec.decode_dev(0, copies[s * pitch: s * pitch + n], n, w, h, dst, stride, mb_rows=(r0, r1))

        ms_mine, ms_max, clk = E.timed(step_stripe, args.steps, warm, sample_clocks=True)
        codec.sync_check(0)
        mine = sha(dst[b0:b1].cpu().numpy())
        untouched = bool(int(dst[:b0].max().item() if b0 > 0 else 0) == 0 and int(dst[b1:].max().item() if b1 < frame_bytes else 0) == 0)
        parts = E.gather_obj((mine == mine_whole, untouched, round(ms_mine, 4)))
        verify["stripe_pixels_sha256_equal_1gpu_decode"] = bool(all(p[0] for p in parts))
        verify["rows_outside_the_stripe_untouched"] = bool(all(p[1] for p in parts))
        extras["stripe_ms_per_rank"] = [p[2] for p in parts]
        extras["independent_frames_per_gpu"] = {
            "mpixel_per_s": round(w * h * E.world * args.steps / (ms_whole_max * 1e-3) / 1e6, 1), "scaling": "weak"}
        step_ms = ms_max / args.steps
        kernel_ms = ms_mine
        # every GPU reads the whole payload (replicated parse); only the pixels are sharded
        alg_bytes = 3 * w * h + n * E.world
    else:
        step_ms = ms_whole_max / args.steps
        kernel_ms = ms_whole
        alg_bytes = 3 * w * h + n
    value = w * h / (step_ms * 1e-3) / 1e6
    launches = args.steps * dec_launches(n, mbw * mbh)

    peak, peak_src = peaks()
    achieved = alg_bytes / (step_ms * 1e-3) / 1e9
    traffic, traffic_src = traffic_from_profile(f"{args.workload}_{args.kind}")
    roofline = {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak * E.world, "unit": "GB/s",
                "frac": round(achieved / (peak * E.world), 4), "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": peak_src + (f" x {E.world} GPUs" if E.world > 1 else ""),
                "kernel": "dec_scan_segments + map tree + dec_emit_offsets + dec_dc_tile x2 + dec_reconstruct "
                          "(one stand-alone decode step; dec_reconstruct is about two thirds of it)",
                "kernel_ms": round(kernel_ms, 4), "algorithmic_bytes_per_launch": alg_bytes,
                "note": "fp64 IDCT in the reference's operation order: issue-bound, not DRAM-bound (DESIGN.md §5)"}

    # ---- end to end: gpu_codec_decode_to_bgr, payload up, pixels down, pinned buffers
    e2e_steps = max(2, min(args.steps, 8))
    e2e = None
    E.barrier()
    if E.rank == 0 or not striped:
        hp, hpay = pinned(lib, n)
        hd, hdst = pinned(lib, frame_bytes)
        torch.from_numpy(hpay).copy_(copies[:n])
        step_whole()
        codec.sync_check(0)
        want = sha(dst.cpu().numpy())
        if striped:
            codec.init(list(range(E.world)))
        try:
            for _ in range(2):
                codec.decode_to_bgr(hpay, w, h, stride, out=hdst)
            if not striped:
                E.barrier()
            t0 = time.perf_counter()
            for _ in range(e2e_steps):
                codec.decode_to_bgr(hpay, w, h, stride, out=hdst)
            dt_ms = (time.perf_counter() - t0) * 1e3
            good = sha(hdst) == want
        finally:
            if striped:
                codec.init([E.local])
            lib.gpu_codec_host_free(C.c_void_p(hp))
            lib.gpu_codec_host_free(C.c_void_p(hd))
    if striped:
        E.host_barrier()
        if E.rank == 0:
            verify["e2e_pixels_sha256_equal_1gpu_decode"] = bool(good)
            e2e = {"value": round(w * h * e2e_steps / (dt_ms * 1e-3) / 1e6, 1), "unit": "Mpixel/s",
                   "h2d_bytes_per_step": n * E.world, "d2h_bytes_per_step": frame_bytes, "steps": e2e_steps,
                   "api": f"gpu_codec_init_devices([0..{E.world - 1}]) + gpu_codec_decode_to_bgr, pinned host buffers, payload "
                          f"to every GPU, a stripe of rows back from each (rank 0 drives all GPUs; wall clock)"}
    else:
        dt_ms = E.max_ms(dt_ms)
        verify["e2e_pixels_sha256_equal_device_resident_decode"] = bool(good)
        e2e = {"value": round(w * h * e2e_steps / (dt_ms * 1e-3) / 1e6, 1), "unit": "Mpixel/s",
               "h2d_bytes_per_step": n, "d2h_bytes_per_step": frame_bytes, "steps": e2e_steps,
               "api": "gpu_codec_decode_to_bgr (pinned host buffers)"}

    out = None
    if E.rank == 0:
        out = {"metric": "decode_mpixel_per_s", "value": round(value, 1), "unit": "Mpixel/s", "n_gpus": E.world,
               "steps": args.steps, "warmup": warm, "ms_per_step": round(step_ms, 4),
               "higher_is_better": True, "scaling": "strong" if striped else "weak", "vs_baseline": None,
               "dtype": DTYPE, "data": "synthetic", "config": make_config(args, E.world), "payload_bytes": n,
               "clocks": clk, "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "verify": verify}
        if E.world == 1:
            out["cpu_baseline"] = cpu_baseline(args)
        if extras:
            out["extras"] = extras
    return out


def run_extras(codec, torch, dev) -> dict:
    """Secondary numbers at N = 1 (never the headline): noise-content 16384x16384 encode."""
    from image_compression_b200 import synth
    ex = {}
    try:
        w = h = 16384
        stride = 3 * w
        rows = torch.empty(h * stride, dtype=torch.uint8, device=dev)
        codec.synth_dev(0, rows, w, h, stride, KIND["noise"], synth.SEED_BASE + 3, 0)
        payload = torch.empty(codec.max_payload_bytes(w, h) + 16, dtype=torch.uint8, device=dev)
        plen = torch.zeros(1, dtype=torch.int64, device=dev)
        for _ in range(3):
            codec.encode_dev(0, rows, w, h, stride, payload, plen)
        codec.sync_check(0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(10):
            codec.encode_dev(0, rows, w, h, stride, payload, plen)
        b.record()
        codec.sync_check(0)
        ms = a.elapsed_time(b) / 10
        n = int(plen.item())
        ex["16384x16384_noise_encode"] = {"payload_bytes": n, "ms_per_frame": round(ms, 4), "mpixel_per_s": round(w * h / ms / 1e3, 1),
                                          "algorithmic_GBps": round((3 * w * h + n) / ms / 1e6, 1)}
    except Exception as e:                      # extras never fail the headline
        ex["error"] = repr(e)
    return ex


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--kind", default="photo", choices=["photo", "noise"])
    ap.add_argument("--no-extras", action="store_true")
    args = ap.parse_args()

    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return 0
        print(json.dumps(run_reference_arm(args)), flush=True)
        return 0
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit(f"bench.py: --gpus {args.gpus} needs one process per GPU: python -m torch.distributed.run "
                         f"--nnodes=1 --nproc-per-node {args.gpus} --master-addr 127.0.0.1 bench.py --gpus {args.gpus} ...")
    E = Env()
    out = run_encode(args, E) if WORKLOADS[args.workload]["op"] == "encode" else run_decode(args, E)
    E.finish()
    if out is not None:
        print(json.dumps(out), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
