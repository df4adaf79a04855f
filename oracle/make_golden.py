#!/usr/bin/env python3
"""Generates tests/golden/*.npz from the UNMODIFIED reference build (oracle/_ref), i.e.
outputs of the reference itself run in this container.  The reference ships no golden
vectors, so these files are what pins the oracle restatement and the CUDA
path on machines where /root/reference does not exist.  Re-run:  python oracle/make_golden.py
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from image_compression_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")

FRAMES = [  # (kind, w, h, seed, frame)
    ("noise", 5, 3, 0x0A11CE01, 0),
    ("photo", 17, 33, 0x0A11CE01, 0),
    ("flat", 48, 16, 0x0A11CE01, 0),
    ("photo", 64, 48, 0x0A11CE02, 1),
    ("noise", 40, 56, 0x0A11CE02, 0),
    ("photo", 250, 130, 0x0A11CE03, 2),
    ("noise", 127, 65, 0x0A11CE03, 0),
]
HASH_FRAMES = [
    ("photo", 640, 427, 0x0A11CE04, 0),  # bigger: only sizes + sha256 kept
    ("photo", 1920, 1080, 0x0A11CE04, 3),
    ("noise", 1024, 512, 0x0A11CE02, 0),
    ("flat", 512, 512, 0x0A11CE02, 0),
]


def sha(b) -> str:
    return hashlib.sha256(bytes(b)).hexdigest()


def main():
    R = oracle.ref()
    assert R is not None, "reference build missing: make -C oracle"
    os.makedirs(OUT, exist_ok=True)
    d = {}
    meta = []
    for i, (kind, w, h, seed, frame) in enumerate(FRAMES):
        rows = synth.frame_bgr_bottomup(kind, w, h, seed, frame)
        awi = R.encode_bmp(oracle.bmp_wrap(rows, w, h))
        bmp = R.decode_to_bmp(awi)
        rgb = R.decode_to_rgb(awi)
        d[f"payload_{i}"] = np.frombuffer(awi, np.uint8)[16:].copy()
        px = np.frombuffer(bmp, np.uint8)[54:].copy().reshape(h, -1)
        px[:, 3 * w:] = 0      # row padding is indeterminate in the reference (un-memset malloc, bmp.c:154)
        d[f"bgr_{i}"] = px.reshape(-1)
        bmp = bmp[:54] + px.tobytes()
        assert rgb.tobytes() == np.frombuffer(bmp, np.uint8)[54:].reshape(h, -1)[::-1, :3 * w].reshape(h, w, 3)[..., ::-1].tobytes()
        meta.append((kind, w, h, seed, frame))
    d["meta"] = np.array([f"{k},{w},{h},{s},{f}" for k, w, h, s, f in meta])
    hashes = []
    for kind, w, h, seed, frame in HASH_FRAMES:
        rows = synth.frame_bgr_bottomup(kind, w, h, seed, frame)
        awi = R.encode_bmp(oracle.bmp_wrap(rows, w, h))
        bmp = R.decode_to_bmp(awi)
        hashes.append(f"{kind},{w},{h},{seed},{frame},{len(awi) - 16},{sha(awi[16:])},{sha(bmp[54:])}")
    d["hashes"] = np.array(hashes)

    # known-answer blocks through the reference's own stage functions
    rng = np.random.default_rng(7)
    yy, xx = np.mgrid[0:8, 0:8]
    blocks = [np.full(64, v, np.uint8) for v in (0, 1, 16, 127, 128, 129, 144, 200, 255)]
    blocks.append((32 * yy + 4 * xx).astype(np.uint8).ravel())
    for u in range(8):
        for v in range(8):
            pat = np.sign(np.cos((2 * yy + 1) * u * np.pi / 16) * np.cos((2 * xx + 1) * v * np.pi / 16))
            blocks.append(np.where(pat > 0, 255, 0).astype(np.uint8).ravel())
            blocks.append(np.where(pat > 0, 0, 255).astype(np.uint8).ravel())
    blocks += [rng.integers(0, 256, 64).astype(np.uint8) for _ in range(200)]
    blocks += [np.clip(rng.normal(128, 6, 64), 0, 255).astype(np.uint8) for _ in range(200)]
    blocks = np.stack(blocks)
    d["kat_blocks"] = blocks
    d["kat_coef"] = np.stack([R.fdct(b) for b in blocks]).astype(np.int16)
    d["kat_chunk_luma"] = np.stack([R.transform(b, 0) for b in blocks]).astype(np.int16)
    d["kat_chunk_chroma"] = np.stack([R.transform(b, 1) for b in blocks]).astype(np.int16)
    d["kat_inv_luma"] = np.stack([R.inverse_transform(c, 0) for c in d["kat_chunk_luma"]])
    d["kat_inv_chroma"] = np.stack([R.inverse_transform(c, 1) for c in d["kat_chunk_chroma"]])
    prevs = rng.integers(-64, 65, len(blocks)).astype(np.int32)
    rle = [R.rle(c, int(p)) for c, p in zip(d["kat_chunk_luma"], prevs)]
    d["kat_rle_prev"] = prevs
    d["kat_rle_len"] = np.array([len(r) for r in rle], np.int32)
    d["kat_rle_bytes"] = np.frombuffer(b"".join(rle), np.uint8)
    # colour: a strided sample of the 2^24 cube + every input where integer division and the
    # reference's fp64 disagree would be caught by the exhaustive tests; keep 8K samples here
    cube = rng.integers(0, 256, (8192, 3)).astype(np.uint8)
    d["colour_in"] = cube
    d["colour_fwd"] = R.rgb2ycc(cube)
    d["colour_inv"] = R.ycc2rgb(cube)
    # whole-cube digests of the reference colour functions
    a = np.arange(256, dtype=np.uint8)
    full = np.stack(np.meshgrid(a, a, a, indexing="ij"), -1).reshape(-1, 3)
    d["colour_cube_sha"] = np.array([sha(R.rgb2ycc(full)), sha(R.ycc2rgb(full))])
    # the reference's sample image: digests only (the image itself stays in /root/reference)
    sample = "/root/reference/sample/sample1.bmp"
    if os.path.exists(sample):
        bmpf = open(sample, "rb").read()
        awi = R.encode_bmp(bmpf)
        d["sample1"] = np.array([str(len(awi)), sha(awi), sha(R.decode_to_bmp(awi)), sha(bmpf)])
        d["sample1_prefix"] = np.frombuffer(awi[:48], np.uint8)
    np.savez_compressed(os.path.join(OUT, "reference_vectors.npz"), **d)
    print("wrote", os.path.join(OUT, "reference_vectors.npz"), os.path.getsize(os.path.join(OUT, "reference_vectors.npz")), "bytes")
    if "sample1" in d:
        print("sample1:", list(d["sample1"]))


if __name__ == "__main__":
    main()
