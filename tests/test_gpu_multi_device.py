"""GPU (-m gpu): one process driving several GPUs through the C ABI (gpu_codec_init(n)): a frame is
split into macroblock-row stripes across the devices and stitched on the host from the prefix sum of
the per-stripe byte lengths; batches go whole-frame per device; decode parses everywhere and
reconstructs a stripe per device.  Skipped on a single-GPU box."""
import numpy as np
import pytest

from image_compression_b200 import synth

pytestmark = pytest.mark.gpu


def test_stripes_and_batches_across_devices(port):
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least two GPUs")
    from image_compression_b200 import codec
    n = min(n, 8)
    codec.init(list(range(n)))
    try:
        for kind, w, h in (("photo", 1000, 700), ("noise", 333, 211), ("photo", 4096, 2048)):
            stride = synth.bmp_stride(w)
            rows = synth.frame_bgr_bottomup(kind, w, h, synth.SEED_BASE + 7)
            want = port.encode_bgr(rows, w, h, stride)
            got = codec.encode_bgr(rows, w, h, stride)                 # n stripes, stitched on the host
            assert got == want.tobytes(), (kind, w, h, n)
            px = codec.decode_to_bgr(got, w, h, stride)                # replicated parse, sharded reconstruct
            assert px.tobytes() == port.decode(want, w, h, stride)[0].tobytes()
            assert codec.decode_to_rgb24(got, w, h).tobytes() == port.decode(want, w, h, rgb_topdown=True)[0].tobytes()
        w, h = 320, 180
        stride = synth.bmp_stride(w)
        frames = [synth.frame_bgr_bottomup("photo", w, h, synth.SEED_BASE + 4, f) for f in range(2 * n + 1)]
        out = codec.encode_batch_bgr(frames, w, h, stride)
        for f, g in zip(frames, out):
            assert g == port.encode_bgr(f, w, h, stride).tobytes()
    finally:
        codec.init([0])        # leave the session's single-device context behind for the other tests
