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
if rank == 0:
    parts = [np.load(os.path.join(tmp, f"seg{r}.npy")).tobytes() for r in range(world)]
    full = oracle.port().encode_bgr(rows, w, h, stride).tobytes()
    assert stripes.stitch(parts) == full
    assert offs[1] == len(parts[0])
    print("STITCH_OK")
dist.destroy_process_group()
