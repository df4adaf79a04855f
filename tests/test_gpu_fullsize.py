"""GPU (-m gpu): BASELINE.json's full sizes.  The oracle is too slow to restate these
whole, so parity rests on (a) the oracle on a band of macroblock rows, (b) size-independent
properties: stripes == whole, discovered chunk offsets == encoder offsets, decode of the
device payload == oracle decode on a band, payload length == last offset + last chunk."""
import numpy as np
import pytest

from image_compression_b200 import stripes, synth

pytestmark = pytest.mark.gpu

KIND = {"noise": 0, "photo": 1, "flat": 2}


def _encode(gpu, rows_t, w, h, stride, mb_rows=None, window=None):
    import torch
    mbw, mbh = (w + 15) // 16, (h + 15) // 16
    r0, r1 = mb_rows or (0, mbh)
    nch = (r1 - r0) * mbw * 6
    payload = torch.empty(nch * 121 + 16, dtype=torch.uint8, device="cuda")
    ln = torch.zeros(1, dtype=torch.int64, device="cuda")
    off = torch.empty(nch, dtype=torch.int32, device="cuda")
    gpu.encode_dev(0, rows_t, w, h, stride, payload, ln, window=window, mb_rows=mb_rows, chunk_off=off)
    gpu.sync_check(0)
    return payload[: int(ln.item())], off


@pytest.mark.parametrize("cfg", [("photo", 3840, 2160, 2), ("noise", 3840, 2160, 2), ("photo", 1920, 1080, 4)])
def test_4k_and_1080p_whole_frame_vs_oracle(gpu, port, cfg):
    import torch
    kind, w, h, cfgno = cfg
    stride = 3 * w
    t = torch.empty(h * stride, dtype=torch.uint8, device="cuda")
    gpu.synth_dev(0, t, w, h, stride, KIND[kind], synth.SEED_BASE + cfgno, 0)
    rows = t.cpu().numpy()
    p, off = _encode(gpu, t, w, h, stride)
    want = port.encode_bgr(rows, w, h, stride)
    assert p.cpu().numpy().tobytes() == want.tobytes()
    dst = torch.zeros(h * stride, dtype=torch.uint8, device="cuda")
    gpu.decode_dev(0, p, p.numel(), w, h, dst, stride)
    gpu.sync_check(0)
    assert dst.cpu().numpy().tobytes() == port.decode(want, w, h, stride)[0].tobytes()


@pytest.mark.parametrize("kind", ["photo", "noise"])
def test_16k_properties(gpu, port, kind):
    """config[2]: 16384x16384.  Oracle on two bands; 8-way stripes == whole; decode round trip."""
    import torch
    w = h = 16384
    stride = 3 * w
    t = torch.empty(h * stride, dtype=torch.uint8, device="cuda")
    gpu.synth_dev(0, t, w, h, stride, KIND[kind], synth.SEED_BASE + 3, 0)
    p, off = _encode(gpu, t, w, h, stride)
    mbw, mbh = w // 16, h // 16
    offs = off.cpu().numpy().view(np.uint32).astype(np.int64)
    assert (np.diff(offs) >= 3).all() and (np.diff(offs) <= 121).all() and offs[0] == 0
    assert 3 <= p.numel() - offs[-1] <= 121
    # oracle on two bands of macroblock rows (the second one starts at a seam)
    for r0, r1 in ((0, 2), (511, 514)):
        y0, y1 = stripes.stripe_window(h, r0, r1)
        b0, b1 = stripes.window_byte_range(h, stride, y0, y1)
        # hand the oracle a virtual full frame: only the window is populated
        win = t[b0:b1].cpu().numpy()
        want = port_encode_window(port, win, w, h, stride, y0, y1, r0, r1)
        lo = offs[r0 * mbw * 6]
        hi = offs[r1 * mbw * 6] if r1 < mbh else p.numel()
        assert p[lo:hi].cpu().numpy().tobytes() == want.tobytes(), (kind, r0)
    # 8-way stripes, each from its own window only
    parts = []
    for r0, r1 in stripes.stripe_rows(mbh, 8):
        y0, y1 = stripes.stripe_window(h, r0, r1)
        b0, b1 = stripes.window_byte_range(h, stride, y0, y1)
        seg, _ = _encode(gpu, t[b0:b1], w, h, stride, mb_rows=(r0, r1), window=(y0, y1))
        parts.append(seg)
    assert torch.equal(torch.cat(parts), p)
    del parts
    # decode: discovery finds the encoder's offsets, pixels agree with side-info decode
    dst = torch.zeros(h * stride, dtype=torch.uint8, device="cuda")
    gpu.decode_dev(0, p, p.numel(), w, h, dst, stride)
    gpu.sync_check(0)
    dst2 = torch.zeros(h * stride, dtype=torch.uint8, device="cuda")
    gpu.decode_dev(0, p, p.numel(), w, h, dst2, stride, chunk_off=off)
    gpu.sync_check(0)
    assert torch.equal(dst, dst2)
    # oracle decode of the first two macroblock rows (predictors start at zero there)
    head = p[: offs[2 * mbw * 6]].cpu().numpy()
    want_px, _ = port.decode(head, w, 32, stride)
    got = dst.view(h, stride)[h - 32:].cpu().numpy()
    assert got.tobytes() == want_px.tobytes()


def port_encode_window(port, win, w, h, stride, y0, y1, r0, r1):
    """Oracle on a window of rows: the oracle addresses rows as rows + (h-1-y)*stride, so give
    it a base pointer shifted as if the whole frame were present (only the window is touched)."""
    import ctypes as C
    lib = port.lib
    nch = (r1 - r0) * ((w + 15) // 16) * 6
    out = np.empty(nch * 121 + 16, np.uint8)
    base = win.ctypes.data - (h - y1) * stride
    lib.awi_oracle_encode_rows.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int,
                                           C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
    n = lib.awi_oracle_encode_rows(C.c_void_p(base), w, h, stride, r0, r1, out.ctypes.data, out.size, None, None)
    return out[:n]


def test_8k_decode_roundtrip(gpu, port):
    """config[4]: 8192x8192 round-trip decode; oracle on the top band, both sinks."""
    import torch
    w = h = 8192
    stride = 3 * w
    t = torch.empty(h * stride, dtype=torch.uint8, device="cuda")
    gpu.synth_dev(0, t, w, h, stride, KIND["photo"], synth.SEED_BASE + 5, 0)
    p, off = _encode(gpu, t, w, h, stride)
    dst = torch.zeros(h * stride, dtype=torch.uint8, device="cuda")
    gpu.decode_dev(0, p, p.numel(), w, h, dst, stride)
    rgb = torch.zeros(h * stride, dtype=torch.uint8, device="cuda")
    gpu.decode_dev(0, p, p.numel(), w, h, rgb, stride, rgb_topdown=True)
    gpu.sync_check(0)
    # the two sinks hold the same picture
    a = dst.view(h, w, 3).flip(0).flip(2)
    assert torch.equal(a, rgb.view(h, w, 3))
    offs = off.cpu().numpy().view(np.uint32).astype(np.int64)
    head = p[: offs[3 * (w // 16) * 6]].cpu().numpy()
    want_px, _ = port.decode(head, w, 48, stride, rgb_topdown=True)
    assert rgb.view(h, stride)[:48].cpu().numpy().tobytes() == want_px.tobytes()
    # lossy but close: PSNR of the round trip stays sane
    err = (dst.view(-1)[:: 97].float() - t.view(-1)[:: 97].float())
    assert float((err * err).mean()) < 120.0


def test_every_colour_through_the_encoder(gpu, port):
    """All 2^24 (R,G,B) triples pass through the encode kernel's own colour rows (dp2a numerators,
    wide-multiply division, exact-multiple fallback): each colour sits once on an even/even pixel,
    so it feeds luma AND both chroma samples.  8192x8192, whole payload against the oracle."""
    import torch
    w = h = 8192
    stride = 3 * w
    idx = torch.arange(1 << 24, dtype=torch.int32, device="cuda").view(4096, 4096)
    rgb = torch.empty((h, w, 3), dtype=torch.uint8, device="cuda")
    for dy, dx, mul in ((0, 0, 1), (0, 1, 65599), (1, 0, 2654435), (1, 1, 40503)):
        c = (idx * mul) & 0xFFFFFF if mul != 1 else idx
        rgb[dy::2, dx::2, 0] = ((c >> 16) & 255).to(torch.uint8)
        rgb[dy::2, dx::2, 1] = ((c >> 8) & 255).to(torch.uint8)
        rgb[dy::2, dx::2, 2] = (c & 255).to(torch.uint8)
    rows = rgb.flip(0).flip(2).contiguous().view(-1)          # bottom-up, B,G,R
    p, _ = _encode(gpu, rows, w, h, stride)
    want = port.encode_bgr(rows.cpu().numpy(), w, h, stride)
    assert p.cpu().numpy().tobytes() == want.tobytes()
