"""Experiment helper: statistics of non-zero AC coefficients per chunk for the bench content, in the thread = chunk
mapping of phase D (warp = 32 consecutive chunks of a tile of 192): mean, mean of per-warp maxima, and the same if the
192 chunks of a tile were sorted by count before being dealt to warps.
  python tools/nz_stats.py [photo|noise] [width height]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from image_compression_b200 import codec, synth
kind = sys.argv[1] if len(sys.argv) > 1 else "photo"
w = int(sys.argv[2]) if len(sys.argv) > 3 else 4096
h = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
codec.init([0])
stride = 3 * w
rows = torch.empty(h * stride, dtype=torch.uint8, device="cuda")
codec.synth_dev(0, rows, w, h, stride, {"noise": 0, "photo": 1}[kind], synth.SEED_BASE + 3, 0)
payload = torch.empty(codec.max_payload_bytes(w, h) + 16, dtype=torch.uint8, device="cuda")
plen = torch.zeros(1, dtype=torch.int64, device="cuda")
n_chunks = (w // 16) * (h // 16) * 6
q = torch.empty(n_chunks * 64, dtype=torch.int8, device="cuda")
codec.encode_dev(0, rows, w, h, stride, payload, plen, qcoef=q)
codec.sync_check(0)
q = q.view(n_chunks, 64)
nz = (q[:, 1:] != 0).sum(1).to(torch.float32)          # AC non-zeros per chunk (what the run-length loop iterates over)
tiles = nz.view(-1, 192)
warps = tiles.view(-1, 6, 32)
print(f"{kind} {w}x{h}: mean nz/chunk {nz.mean():.2f}, max {nz.max():.0f}")
print(f"  per-warp max (tid = chunk):   mean {warps.max(2).values.mean():.2f}   sum per tile {warps.max(2).values.sum(1).mean():.1f}   CTA max {tiles.max(1).values.mean():.2f}")
s = tiles.sort(1).values.view(-1, 6, 32)
print(f"  per-warp max (sorted in tile): mean {s.max(2).values.mean():.2f}   sum per tile {s.max(2).values.sum(1).mean():.1f}")
# interleaved sort: warp k gets ranks k, k+6, ... (balances warps; same per-warp max as CTA max mostly)
# bucketed by 4 popcount classes
for nb in (2, 4, 8):
    edges = torch.quantile(nz, torch.linspace(0, 1, nb + 1, device=nz.device)[1:-1])
    cls = torch.bucketize(tiles, edges)
    order = torch.argsort(cls, dim=1, stable=True)
    s = torch.gather(tiles, 1, order).view(-1, 6, 32)
    print(f"  per-warp max ({nb} count classes): mean {s.max(2).values.mean():.2f}   sum per tile {s.max(2).values.sum(1).mean():.1f}")
