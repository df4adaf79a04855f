#!/usr/bin/env python3
"""Experiments only: build a variant of libgpu_codec.so with extra nvcc flags into
image_compression_b200/variants/libgpu_codec_<name>.so (git-ignored like every .so; it travels to the GPU box).
Use it with  GPU_CODEC_LIB=<path> python bench.py ...  — the product path always loads the in-tree libgpu_codec.so.

  python tools/build_variant.py cta3 -DENC_MIN_CTAS=3
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from image_compression_b200 import build as B  # noqa: E402


def main():
    name, extra = sys.argv[1], sys.argv[2:]
    out_dir = os.path.join(B.HERE, "variants")
    tmp = os.path.join(out_dir, "obj_" + name)
    os.makedirs(tmp, exist_ok=True)
    objs, procs = [], []
    for src in B.SOURCES:
        obj = os.path.join(tmp, src.replace(".cu", ".o"))
        cmd = ["nvcc", *B.NVCC_FLAGS, "-Xptxas=-v", *extra, "-c", os.path.join(B.CSRC, src), "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            sys.stderr.write(out)
            raise SystemExit(f"nvcc failed on {src}")
        if src == "awi_encode.cu":
            lines = out.splitlines()
            for i, ln in enumerate(lines):
                if "awi_encode_kernelENS" in ln and "Compiling" in ln:
                    print(name, "|", " ".join(x.strip() for x in lines[i + 2:i + 4]))
    lib = os.path.join(out_dir, f"libgpu_codec_{name}.so")
    subprocess.run(["nvcc", "-shared", "-o", lib, *objs, "-lcudart"], check=True, stderr=subprocess.DEVNULL)
    print(lib)


if __name__ == "__main__":
    main()
