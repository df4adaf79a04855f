"""Helper for test_host_logic.py::test_two_rank_stripe_stitch_gloo (not collected by pytest)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from image_compression_b200 import stripes, synth  # noqa: E402

rank, world, tmp = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
dist.init_process_group("gloo", rank=rank, world_size=world)
w, h = 208, 176
stride = synth.bmp_stride(w)
rows = synth.frame_bgr_bottomup("photo", w, h, 0xA11CE03).reshape(-1)
mbh = (h + 15) // 16
r0, r1 = stripes.stripe_rows(mbh, world)[rank]
y0, y1 = stripes.stripe_window(h, r0, r1)
b0, b1 = stripes.window_byte_range(h, stride, y0, y1)
# the rank only looks at its window: blank everything else to prove the halo is sufficient
mine = np.zeros_like(rows)
mine[b0:b1] = rows[b0:b1]
seg = oracle.port().encode_rows(mine, w, h, stride, r0, r1)
lens = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
dist.all_gather(lens, torch.tensor([seg.size], dtype=torch.int64))
offs = stripes.segment_offsets([int(x) for x in lens])
np.save(os.path.join(tmp, f"seg{rank}.npy"), seg)
dist.barrier()
# what bench.py does at N > 1: every rank publishes (length, sha256) of its segment, and the segments are checked
# against slices of the single-device payload at the prefix-sum offsets
import hashlib  # noqa: E402
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _bands  # noqa: E402
full, full_lens = oracle.port().encode_bgr(rows, w, h, stride, want_coefs=True)[::2]
parts = [None] * world
dist.all_gather_object(parts, (int(seg.size), hashlib.sha256(seg.tobytes()).hexdigest()))
offs2 = stripes.segment_offsets([p[0] for p in parts])
assert offs2 == offs
assert all(hashlib.sha256(full[o:o + n].tobytes()).hexdigest() == d for o, (n, d) in zip(offs2, parts))
# decode side: replicated parse (every rank sees the whole payload and derives offsets + predictors itself),
# sharded reconstruct (only rows [r0, r1) are written), no exchange
mbw = (w + 15) // 16
coff = np.concatenate([[0], np.cumsum(full_lens.astype(np.int64))])
dc = _bands.dc_predictors(full, coff[:-1])
dst = np.zeros(h * stride, np.uint8)
pred = dc[r0 * mbw - 1] if r0 > 0 else np.zeros(6, np.int32)
end = oracle.port().decode_rows(full, w, h, r0, r1, int(coff[r0 * mbw * 6]), pred, dst, stride)
assert end == int(coff[r1 * mbw * 6])
np.save(os.path.join(tmp, f"px{rank}.npy"), dst)
dist.barrier()
if rank == 0:
    parts = [np.load(os.path.join(tmp, f"seg{r}.npy")).tobytes() for r in range(world)]
    assert stripes.stitch(parts) == full.tobytes()
    assert offs[1] == len(parts[0])
    px = sum(np.load(os.path.join(tmp, f"px{r}.npy")).astype(np.int32) for r in range(world)).astype(np.uint8)   # disjoint rows
    assert px.tobytes() == oracle.port().decode(full, w, h, stride)[0].tobytes()
    print("STITCH_OK")
dist.destroy_process_group()
