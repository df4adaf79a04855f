"""GPU (-m gpu): BASELINE.json's full sizes, whole frames against the oracle.  One oracle thread would
take minutes on 16384x16384, so the frame is cut into bands of macroblock rows that run on all host
threads (tests/_bands.py; the bands are exact because only six DC predictors cross a band seam); on top
of that the size-independent properties: stripes == whole, discovered chunk offsets == encoder offsets."""
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
def test_16k_whole_frame_vs_oracle(gpu, port, ref, kind):
    """config[2]: 16384x16384, EVERY byte against the oracle (bands of macroblock rows on the host's threads,
    tests/_bands.py): payload, chunk offsets, decoded pixels; photo-like content also against the unmodified
    reference build where it travelled.  Plus the size-independent properties: 8-way stripes == whole,
    discovered chunk offsets == encoder offsets."""
    import torch
    import oracle
    import _bands
    w = h = 16384
    stride = 3 * w
    t = torch.empty(h * stride, dtype=torch.uint8, device="cuda")
    gpu.synth_dev(0, t, w, h, stride, KIND[kind], synth.SEED_BASE + 3, 0)
    p, off = _encode(gpu, t, w, h, stride)
    mbw, mbh = w // 16, h // 16
    offs = off.cpu().numpy().view(np.uint32).astype(np.int64)
    rows = t.cpu().numpy()
    got = p.cpu().numpy()
    # whole frame, restatement: bytes and chunk boundaries
    want, lens = _bands.port_encode_frame(port, rows, w, h, stride)
    assert got.size == want.size and got.tobytes() == want.tobytes(), kind
    assert (offs == np.concatenate([[0], np.cumsum(lens.astype(np.int64))])[:-1]).all()
    # whole frame, the reference's own compress_to_awi (the unmodified sources)
    if ref is not None and kind == "photo":
        _bands.ref_check_frame(ref, oracle, rows, w, h, stride, got, offs)
    del rows
    # 8-way stripes, each from its own window only
    parts = []
    for r0, r1 in stripes.stripe_rows(mbh, 8):
        y0, y1 = stripes.stripe_window(h, r0, r1)
        b0, b1 = stripes.window_byte_range(h, stride, y0, y1)
        seg, _ = _encode(gpu, t[b0:b1], w, h, stride, mb_rows=(r0, r1), window=(y0, y1))
        parts.append(seg)
    assert torch.equal(torch.cat(parts), p)
    del parts
    # decode: stand-alone (boundary discovery) and with the encoder's offsets, both against the oracle's pixels
    want_px = _bands.port_decode_frame(port, want, lens, w, h, stride)
    dst = torch.zeros(h * stride, dtype=torch.uint8, device="cuda")
    gpu.decode_dev(0, p, p.numel(), w, h, dst, stride)
    gpu.sync_check(0)
    assert dst.cpu().numpy().tobytes() == want_px.tobytes(), kind
    dst.zero_()
    gpu.decode_dev(0, p, p.numel(), w, h, dst, stride, chunk_off=off)
    gpu.sync_check(0)
    assert dst.cpu().numpy().tobytes() == want_px.tobytes(), kind


@pytest.mark.parametrize("kind", ["photo", "noise"])
def test_8k_decode_whole_frame_vs_oracle(gpu, port, ref, kind):
    """config[4]: 8192x8192 round-trip decode, every pixel of both sinks against the oracle (and the encoder's
    payload against it too, since that is what gets decoded)."""
    import torch
    import oracle
    import _bands
    w = h = 8192
    stride = 3 * w
    t = torch.empty(h * stride, dtype=torch.uint8, device="cuda")
    gpu.synth_dev(0, t, w, h, stride, KIND[kind], synth.SEED_BASE + 5, 0)
    p, off = _encode(gpu, t, w, h, stride)
    rows = t.cpu().numpy()
    want, lens = _bands.port_encode_frame(port, rows, w, h, stride)
    got = p.cpu().numpy()
    assert got.tobytes() == want.tobytes()
    if ref is not None:
        offs = off.cpu().numpy().view(np.uint32).astype(np.int64)
        _bands.ref_check_frame(ref, oracle, rows, w, h, stride, got, offs)
    dst = torch.zeros(h * stride, dtype=torch.uint8, device="cuda")
    gpu.decode_dev(0, p, p.numel(), w, h, dst, stride)
    rgb = torch.zeros(h * stride, dtype=torch.uint8, device="cuda")
    gpu.decode_dev(0, p, p.numel(), w, h, rgb, stride, rgb_topdown=True)
    gpu.sync_check(0)
    assert dst.cpu().numpy().tobytes() == _bands.port_decode_frame(port, want, lens, w, h, stride).tobytes()
    assert rgb.cpu().numpy().tobytes() == _bands.port_decode_frame(port, want, lens, w, h, stride, True).tobytes()
    # through the host entry points too (upload, sub-stripe pipeline, download)
    assert gpu.decode_to_bgr(got, w, h, stride).tobytes() == dst.cpu().numpy().tobytes()
    if kind == "photo":
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


def test_long_batch_takes_the_three_kernel_path(gpu, port):
    """Above 65536 tiles the tile scan and the gather are kernels of their own again (enc_scan_tiles + enc_gather_tiles):
    260 x 1080p frames (66 300 tiles, 1.6 GB of pixels) in one gpu_codec_encode_batch_bgr_dev call.  Every frame's payload must
    equal what the same frame gives on its own (through the one-kernel gather), and the oracle's for a sample of them; a batch
    of 200 frames (51 000 tiles: per-64-tile sums inside one kernel) is checked the same way."""
    import torch
    w, h = 1920, 1080
    stride = 3 * w
    fb = stride * h
    n_max = 260
    buf = torch.empty(n_max * fb, dtype=torch.uint8, device="cuda")
    for f in range(n_max):
        gpu.synth_dev(0, buf[f * fb:(f + 1) * fb], w, h, stride, KIND["photo" if f % 3 else "noise"], synth.SEED_BASE + 4, f)
    cap = gpu.max_payload_bytes(w, h) + 16
    single = torch.empty(cap, dtype=torch.uint8, device="cuda")
    ln = torch.zeros(1, dtype=torch.int64, device="cuda")
    ref_digest = {}
    for f in (0, 1, 63, 64, 127, 199, 200, 258, 259):
        gpu.encode_dev(0, buf[f * fb:(f + 1) * fb], w, h, stride, single, ln)
        gpu.sync_check(0)
        ref_digest[f] = single[: int(ln.item())].cpu().numpy().tobytes()
    assert ref_digest[0] == port.encode_bgr(buf[:fb].cpu().numpy(), w, h, stride).tobytes()
    assert ref_digest[259] == port.encode_bgr(buf[259 * fb:260 * fb].cpu().numpy(), w, h, stride).tobytes()
    payload = torch.empty(n_max * 1200000, dtype=torch.uint8, device="cuda")     # photo frames take ~0.36 MB, noise ~2 MB: 235 MB in all
    for n in (260, 200):
        off = torch.zeros(n + 1, dtype=torch.int64, device="cuda")
        for _ in range(2):                                                    # twice: the second call meets the first one's leftovers
            gpu.encode_batch_dev(0, buf, n, fb, w, h, stride, payload, off)
            gpu.sync_check(0)
        o = off.cpu().numpy()
        assert o[0] == 0 and (np.diff(o) > 0).all()
        for f, want in ref_digest.items():
            if f < n:
                assert payload[o[f]:o[f + 1]].cpu().numpy().tobytes() == want, (n, f)
