"""hogb200_block_histograms_noalloc (no counterpart in the reference): the
fused gradient + histogram kernel with its scratch in a static __device__ array, through the C ABI, against the existing
hogb200_block_histograms on two calls in a row (the scratch keeps the previous call's contents).  Collected last on
purpose."""
import numpy as np
import pytest

from _oracle import synthetic_frame

pytestmark = pytest.mark.gpu

DESC_TOL = 1e-4


def test_block_histograms_noalloc_matches_existing_entry_point():
    import torch
    from opencv_b200 import HOG, HogError, load_library
    assert load_library().hogb200_noalloc_scratch_bytes() == 256 * 1024
    h = HOG.create()
    # two calls in a row, different shapes and channel counts: the second one finds the first one's scratch contents
    for seed, (w, hh, cn) in enumerate([(640, 480, 4), (333, 222, 1), (1920, 1080, 4), (640, 480, 4)]):
        img = synthetic_frame(50 + seed, w, hh, cn)
        want = h.blockHistograms(img)
        pitch = (w * cn + 15) // 16 * 16                       # rows 16-byte aligned, as the entry point demands
        buf = torch.zeros((hh, pitch), dtype=torch.uint8, device="cuda")
        buf[:, :w * cn] = torch.from_numpy(img.reshape(hh, w * cn)).cuda()
        view = buf[:, :w * cn].view(hh, w, cn) if cn == 4 else buf[:, :w]
        out = torch.full(want.shape, -1.0, dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        for rep in range(2):
            out.fill_(-1.0)
            torch.cuda.synchronize()
            bx, by = h.blockHistogramsNoAlloc(view, out)
            torch.cuda.synchronize()
            assert (by, bx) == want.shape[:2]
            got = out.cpu().numpy()
            assert np.abs(got - want).max() <= DESC_TOL, (w, hh, cn, rep, float(np.abs(got - want).max()))
        assert torch.cuda.mem_get_info()[0] >= free0           # the calls allocated no device memory
    # host images, unaligned rows and images beyond the static scratch are refused with a clear error
    out = torch.zeros((10,), dtype=torch.float32, device="cuda")
    with pytest.raises(HogError) as e:
        h.blockHistogramsNoAlloc(synthetic_frame(1, 64, 64, 1), out)
    assert e.value.status == 1 and "device memory" in str(e.value)
    odd = torch.zeros((64, 67), dtype=torch.uint8, device="cuda")
    with pytest.raises(HogError) as e:
        h.blockHistogramsNoAlloc(odd[:, :66], out)
    assert e.value.status == 1
    big = torch.zeros((20000, 30000), dtype=torch.uint8, device="cuda")      # 2499 x 3749 blocks: 42 076 work items > 32 704
    cap = torch.zeros((2499 * 3749 * 36,), dtype=torch.float32, device="cuda")
    with pytest.raises(HogError) as e:
        h.blockHistogramsNoAlloc(big, cap)
    assert e.value.status == 3 and "static scratch" in str(e.value)
