#!/usr/bin/env python3
"""bench.py — encode Mpixel/s of the AWI hot path on B200, one JSON line on stdout.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                  [--workload c3|c2|c4] [--kind photo|noise]

A "step" is one pass of the fused encode kernel over one frame resident in HBM.
Default workload c3 = BASELINE.json configs[2]: a synthetic 16384x16384 24-bit frame
(805 MB, larger than L2, so no flush is needed between iterations).  With N ranks (one
process per GPU, torchrun) every rank encodes its own such frame — independent units, no
data-path collective, weak scaling; NCCL is used only for the barrier and the max-over-ranks
of the device-side time.

`value`     device-resident throughput (CUDA events on the launching stream).
`e2e`       same metric through the reference-facing C ABI call gpu_codec_encode_bgr with
            pinned HOST buffers: H2D of the frame and D2H of the payload inside the timed region.
`roofline`  algorithmic bytes (3WH read + payload written) / kernel time vs the measured HBM copy peak.
`cpu_baseline` the reference's own C code (oracle/_ref, single thread — the reference has no
            threads) on a bounded band of the same frame, timed on this box's host.
`--impl reference` times that reference code on all host cores (one band per thread).
`extras`    secondary numbers, never the headline: at N = 1 noise-content encode and the 8192x8192 decode
            (device-resident and end to end); at N > 1 ONE 16384x16384 frame cut into N macroblock-row stripes,
            a stripe per GPU (configs[2] read literally, strong scaling) with the check that the stripes' byte
            lengths add up to the whole frame's payload.
"""
from __future__ import annotations

import argparse
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
    # name: (width, height, frames per rank, BASELINE.json config it restates)
    "c3": (16384, 16384, 1, "configs[2]: synthetic 16384x16384 BGR frame encode (one frame per GPU)"),
    "c2": (3840, 2160, 1, "configs[1]: synthetic 3840x2160 BGR single-frame encode"),
    "c4": (1920, 1080, 128, "configs[3]: batch of 1920x1080 frames, 128 per GPU, image-parallel"),
}
KIND = {"noise": 0, "photo": 1, "flat": 2}
CFG_NO = {"c2": 2, "c3": 3, "c4": 4}
L2_BYTES = 126 << 20


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


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


# ----------------------------------------------------------------------------- reference arm / CPU baseline

def _reference_lib():
    import oracle
    r = oracle.ref()
    if r is not None:
        return r, "reference"
    return oracle.port(), "port"


def _band_bmp(width: int, band_rows: int, kind: str, seed: int) -> np.ndarray:
    """A BMP file image of the top `band_rows` rows of the workload frame (same generator,
    same seed: the first rows of the frame are the first rows of the band)."""
    import oracle
    from image_compression_b200 import synth
    rows = synth.frame_bgr_bottomup(kind, width, band_rows, seed, 0)
    return np.frombuffer(oracle.bmp_wrap(rows, width, band_rows), np.uint8)


def cpu_encode_once(lib, kind_name, bmp: np.ndarray, width: int, rows: int) -> float:
    t = time.perf_counter()
    if kind_name == "reference":
        lib.encode_bmp(bmp)
    else:
        lib.encode_bgr(bmp[54:], width, rows, 3 * width)
    return time.perf_counter() - t


def cpu_baseline(width: int, height: int, kind: str, seed: int) -> dict:
    lib, kind_name = _reference_lib()
    band = min(height, 2048 if width >= 8192 else height)
    bmp = _band_bmp(width, band, kind, seed)
    dt = min(cpu_encode_once(lib, kind_name, bmp, width, band) for _ in range(2))
    return {"value": round(width * band / dt / 1e6, 3), "unit": "Mpixel/s", "cores": 1, "kind": kind_name,
            "sample": f"top {band} rows of the {width}x{height} {kind} frame ({width * band / 1e6:.1f} Mpx), "
                      f"compress_to_awi from memory, gcc -O3 -ffp-contract=off, best of 2",
            "seconds": round(dt, 3)}


def run_reference_arm(args, width, height, workload_desc) -> dict:
    lib, kind_name = _reference_lib()
    cores = os.cpu_count() or 1
    band = 256 if width >= 8192 else min(height, 1088)
    seed = _seed(args.workload)
    bmp = _band_bmp(width, band, args.kind, seed)
    px_per_step = width * band * cores

    def step():
        ts = [threading.Thread(target=cpu_encode_once, args=(lib, kind_name, bmp, width, band)) for _ in range(cores)]
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
    sample = (f"{cores} threads x a {width}x{band} band of the {width}x{height} {args.kind} frame per step "
              f"({px_per_step / 1e6:.1f} Mpx/step); the reference's compress_to_awi, memory-backed")
    return {"impl": "reference", "metric": "encode_mpixel_per_s", "value": val, "unit": "Mpixel/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32+f64 (bit-exact u8 in/out)",
            "data": "synthetic", "config": {"workload": workload_desc, "kind": args.kind, "width": width, "height": height},
            "cpu_baseline": {"value": val, "unit": "Mpixel/s", "cores": cores, "kind": kind_name, "sample": sample},
            "e2e": {"value": val, "unit": "Mpixel/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


def _seed(workload: str) -> int:
    from image_compression_b200 import synth
    return synth.SEED_BASE + CFG_NO[workload]


# ----------------------------------------------------------------------------- our arm

def run_ours(args, width, height, frames_per_rank, workload_desc) -> dict | None:
    import torch
    import torch.distributed as dist
    from image_compression_b200 import codec

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the codec has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    codec.init([local])
    dev = torch.device("cuda", local)
    stride = 3 * width
    frame_bytes = stride * height
    seed = _seed(args.workload)
    kind = KIND[args.kind]

    # resident inputs: frames_per_rank distinct frames (frame index unique across ranks)
    n_res = frames_per_rank
    rows = torch.empty((n_res, frame_bytes), dtype=torch.uint8, device=dev)
    for f in range(n_res):
        codec.synth_dev(0, rows[f], width, height, stride, kind, seed, rank * frames_per_rank + f)
    cap = codec.max_payload_bytes(width, height) + 16
    payload = torch.empty(cap, dtype=torch.uint8, device=dev)
    plen = torch.zeros(1, dtype=torch.int64, device=dev)
    codec.sync_check(0)

    batch_off = torch.zeros(n_res + 1, dtype=torch.int64, device=dev)
    if n_res > 1:
        payload = torch.empty(cap * n_res, dtype=torch.uint8, device=dev)

    def step():
        if n_res == 1:
            codec.encode_dev(0, rows[0], width, height, stride, payload, plen)
        else:      # image-parallel batch: one set of launches for all resident frames
            codec.encode_batch_dev(0, rows, n_res, frame_bytes, width, height, stride, payload, batch_off)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    codec.sync_check(0)
    payload_bytes = int(plen.item()) if n_res == 1 else int(batch_off[-1].item())

    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    with ClockSampler(local) as clk:
        ev0.record()
        for _ in range(args.steps):
            step()
        ev1.record()
        barrier()
    codec.sync_check(0)
    ms = ev0.elapsed_time(ev1)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    px_per_step = width * height * frames_per_rank * world
    value = px_per_step * args.steps / (ms_max * 1e-3) / 1e6

    # roofline of the fused encode kernel (the only kernel in the step)
    peak, peak_src = peaks()
    # awi_encode_kernel + enc_scan_tiles + enc_gather_tiles per step; up to 4096 tiles the last two are one kernel
    n_tiles = ((((width + 15) // 16) * ((height + 15) // 16)) + 31) // 32 * frames_per_rank
    launches = args.steps * (2 if n_tiles <= 4096 else 3)
    alg_bytes = 3 * width * height * frames_per_rank + payload_bytes      # per step
    kernel_ms = ms / args.steps   # one encode step = awi_encode_kernel + tile scan + gather (the first is >95 %)
    achieved = alg_bytes / (kernel_ms * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get(f"{args.workload}_{args.kind}")
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                "frac": round(achieved / peak, 4), "traffic": traffic, "peak_source": peak_src,
                "kernel": "awi_encode_kernel (+ enc_scan_tiles, enc_gather_tiles)", "kernel_ms": round(kernel_ms, 4), "algorithmic_bytes_per_launch": alg_bytes,
                "note": "integer/fp32 issue-bound under the bit-exactness rule, not DRAM-bound (DESIGN.md)"}

    # end to end through the C ABI with host buffers (pinned), H2D + D2H inside the timed region
    e2e_steps = max(2, min(args.steps, 8))
    lib = codec.load()
    hptr = lib.gpu_codec_host_alloc(frame_bytes)
    if not hptr:
        raise SystemExit("bench.py: pinned allocation failed")
    import ctypes as C
    host = np.ctypeslib.as_array(C.cast(hptr, C.POINTER(C.c_uint8)), shape=(frame_bytes,))
    torch.from_numpy(host).copy_(rows[0])
    torch.cuda.synchronize()
    e2e_payload = 0
    if frames_per_rank > 1:        # image-parallel batch through gpu_codec_encode_batch_bgr: all frames from pinned memory
        hbatch = lib.gpu_codec_host_alloc(frame_bytes * frames_per_rank)
        hb = np.ctypeslib.as_array(C.cast(hbatch, C.POINTER(C.c_uint8)), shape=(frames_per_rank, frame_bytes))
        torch.from_numpy(hb).copy_(rows)
        torch.cuda.synchronize()
        ptrs = (C.c_void_p * frames_per_rank)(*[hbatch + f * frame_bytes for f in range(frames_per_rank)])
        outs = (C.c_void_p * frames_per_rank)()
        lens = (C.c_size_t * frames_per_rank)()

        def e2e_step():
            if not lib.gpu_codec_encode_batch_bgr(ptrs, frames_per_rank, width, height, stride, outs, lens):
                raise SystemExit("bench.py: " + lib.gpu_codec_last_error().decode())
            tot = sum(lens[i] for i in range(frames_per_rank))
            for i in range(frames_per_rank):
                lib.gpu_codec_free(C.c_void_p(outs[i]))
            return tot
    else:
        def e2e_step():
            addr, n = codec.encode_bgr_ptr(hptr, width, height, stride)
            codec.free_payload(addr)
            return n
    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    ev0.record()
    for _ in range(e2e_steps):
        e2e_payload = e2e_step()
    ev1.record()
    torch.cuda.synchronize()
    e2e_ms = max(ev0.elapsed_time(ev1), (time.perf_counter() - t0) * 1e3)
    te = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_val = px_per_step * e2e_steps / (float(te.item()) * 1e-3) / 1e6
    lib.gpu_codec_host_free(C.c_void_p(hptr))
    e2e = {"value": round(e2e_val, 1), "unit": "Mpixel/s", "h2d_bytes_per_step": frame_bytes * frames_per_rank,
           "d2h_bytes_per_step": e2e_payload + 8 * (frames_per_rank + 1), "steps": e2e_steps,
           "api": ("gpu_codec_encode_bgr" if frames_per_rank == 1 else "gpu_codec_encode_batch_bgr") + " (pinned host buffers)"}

    extras = {}
    if rank == 0 and world == 1 and not args.no_extras:
        extras = run_extras(codec, torch, dev, args)
    if world > 1 and frames_per_rank == 1 and not args.no_extras:
        # configs[2] read literally: ONE frame (rank 0's), cut into macroblock-row stripes, a stripe per GPU
        # (strong scaling; every rank holds the frame and encodes rows [r0, r1) with their halo row).  The
        # stripes' byte lengths must add up to the whole frame's payload length.
        codec.synth_dev(0, rows[0], width, height, stride, kind, seed, 0)
        mbh = (height + 15) // 16
        r0, r1 = mbh * rank // world, mbh * (rank + 1) // world
        slen = torch.zeros(1, dtype=torch.int64, device=dev)
        for _ in range(3):
            codec.encode_dev(0, rows[0], width, height, stride, payload, slen, mb_rows=(r0, r1))
        barrier()
        ev0.record()
        for _ in range(args.steps):
            codec.encode_dev(0, rows[0], width, height, stride, payload, slen, mb_rows=(r0, r1))
        ev1.record()
        barrier()
        codec.sync_check(0)
        ts = torch.tensor([ev0.elapsed_time(ev1) / args.steps], dtype=torch.float64, device=dev)
        tot = slen.clone()
        dist.all_reduce(ts, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        whole = torch.tensor([payload_bytes if rank == 0 else 0], dtype=torch.int64, device=dev)
        dist.all_reduce(whole, op=dist.ReduceOp.SUM)
        extras = {"one_frame_in_stripes": {
            "ms_per_frame": round(float(ts.item()), 4),
            "mpixel_per_s": round(width * height / (float(ts.item()) * 1e-3) / 1e6, 1),
            "scaling": "strong", "stripes": world,
            "stripe_bytes_sum_equals_frame_payload": bool(int(tot.item()) == int(whole.item()))}}

    out = None
    if rank == 0:
        out = {"metric": "encode_mpixel_per_s", "value": round(value, 1), "unit": "Mpixel/s", "n_gpus": world,
               "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": round(ms_max / args.steps, 4),
               "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
               "dtype": "f32+f64 (bit-exact u8 in/out)", "data": "synthetic",
               "config": {"workload": workload_desc, "kind": args.kind, "width": width, "height": height,
                          "frames_per_gpu": frames_per_rank, "payload_bytes": payload_bytes,
                          "l2": "inputs larger than L2 (%.0f MB per step vs 126 MB), no flush needed"
                                % (frame_bytes * frames_per_rank / 1e6) if frame_bytes * frames_per_rank > 2 * L2_BYTES
                          else "inputs fit L2: rotate more frames for a cold-cache number",
                          "parallelism": f"{world} x independent frames, no collective"},
               "clocks": clk.summary(), "e2e": e2e, "gpu_launches": launches, "roofline": roofline}
        if world == 1:
            out["cpu_baseline"] = cpu_baseline(width, height, args.kind, seed)
        if extras:
            out["extras"] = extras
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out


def run_extras(codec, torch, dev, args) -> dict:
    """Secondary numbers (not the headline): noise-content encode and configs[4] decode."""
    from image_compression_b200 import synth
    ex = {}

    def timed(fn, steps):
        for _ in range(3):
            fn()
        codec.sync_check(0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            fn()
        b.record()
        codec.sync_check(0)
        return a.elapsed_time(b) / steps

    try:
        w = h = 8192
        stride = 3 * w
        for kind in ("photo", "noise"):
            rows = torch.empty(h * stride, dtype=torch.uint8, device=dev)
            codec.synth_dev(0, rows, w, h, stride, KIND[kind], synth.SEED_BASE + 5, 0)
            payload = torch.empty(codec.max_payload_bytes(w, h) + 16, dtype=torch.uint8, device=dev)
            plen = torch.zeros(1, dtype=torch.int64, device=dev)
            off = torch.empty((w // 16) * (h // 16) * 6, dtype=torch.int32, device=dev)
            ms_e = timed(lambda: codec.encode_dev(0, rows, w, h, stride, payload, plen, chunk_off=off), 10)
            n = int(plen.item())
            dst = torch.empty(h * stride, dtype=torch.uint8, device=dev)
            ms_d = timed(lambda: codec.decode_dev(0, payload, n, w, h, dst, stride), 10)
            ms_s = timed(lambda: codec.decode_dev(0, payload, n, w, h, dst, stride, chunk_off=off), 10)
            ex[f"8192x8192_{kind}"] = {
                "payload_bytes": n, "encode_mpixel_per_s": round(w * h / ms_e / 1e3, 1),
                "decode_standalone_mpixel_per_s": round(w * h / ms_d / 1e3, 1),
                "decode_with_encoder_offsets_mpixel_per_s": round(w * h / ms_s / 1e3, 1),
                "decode_algorithmic_GBps": round((3 * w * h + n) / ms_d / 1e6, 1)}
            if kind == "photo":
                # the same decode end to end through gpu_codec_decode_to_bgr: payload up, pixels down (pinned buffers)
                import ctypes as C
                import numpy as np
                lib = codec.load()
                hp, hd = lib.gpu_codec_host_alloc(n), lib.gpu_codec_host_alloc(h * stride)
                if hp and hd:
                    hpay = np.ctypeslib.as_array(C.cast(hp, C.POINTER(C.c_uint8)), shape=(n,))
                    hdst = np.ctypeslib.as_array(C.cast(hd, C.POINTER(C.c_uint8)), shape=(h * stride,))
                    torch.from_numpy(hpay).copy_(payload[:n])
                    torch.cuda.synchronize()
                    codec.decode_to_bgr(hpay, w, h, stride, out=hdst)
                    t0 = time.perf_counter()
                    for _ in range(5):
                        codec.decode_to_bgr(hpay, w, h, stride, out=hdst)
                    ex[f"8192x8192_{kind}"]["decode_e2e_mpixel_per_s"] = round(w * h * 5 / (time.perf_counter() - t0) / 1e6, 1)
                    ok = bool((torch.from_numpy(hdst).to(dev) == dst).all().item())
                    ex[f"8192x8192_{kind}"]["decode_e2e_matches_device_path"] = ok
                for ptr in (hp, hd):
                    if ptr:
                        lib.gpu_codec_host_free(C.c_void_p(ptr))
            del rows, payload, dst
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
    width, height, frames, desc = WORKLOADS[args.workload]

    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return 0
        print(json.dumps(run_reference_arm(args, width, height, desc)), flush=True)
        return 0
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit(f"bench.py: --gpus {args.gpus} needs one process per GPU: python -m torch.distributed.run "
                             f"--nnodes=1 --nproc-per-node {args.gpus} --master-addr 127.0.0.1 bench.py --gpus {args.gpus} ...")
    out = run_ours(args, width, height, frames, desc)
    if out is not None:
        print(json.dumps(out), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
