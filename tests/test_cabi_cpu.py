"""CPU-side checks of the product library: it loads, exports every symbol include/hogb200.h declares, fails loudly
without a GPU, and its host-only logic (level enumeration, groupRectangles) equals the oracle's."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from _oracle import Port, Ref
from opencv_b200 import HOG, HogError, load_library, lib_path

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    L = C.CDLL(lib_path())
    for header, prefix, at_least in (("hogb200.h", "hogb200_", 38), ("tbdb200.h", "tbdb200_", 14)):
        hdr = open(os.path.join(ROOT, "include", header)).read()
        hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
        names = set(re.findall(r"\b(%s[a-z0-9_]+)\s*\(" % prefix, hdr))
        assert len(names) >= at_least, (header, len(names))
        missing = [n for n in sorted(names) if not hasattr(L, n)]
        assert not missing, missing


def test_create_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(HogError) as e:
        HOG.create()
    assert e.value.status == 2 and "no CPU fallback" in str(e.value)


def test_default_people_detector_tables():
    L = load_library()
    for (w, h, n) in [(64, 128, 3781), (48, 96, 1981)]:
        p, cnt = C.c_void_p(), C.c_size_t()
        assert L.hogb200_default_people_detector(w, h, C.byref(p), C.byref(cnt)) == 0
        assert cnt.value == n
        d = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(n,))
        if Ref.available():
            assert np.array_equal(d, Ref().people_detector(0 if w == 64 else 1))
    p, cnt = C.c_void_p(), C.c_size_t()
    assert L.hogb200_default_people_detector(32, 64, C.byref(p), C.byref(cnt)) == 1     # CV_Assert in the reference


@pytest.mark.parametrize("shape", [(640, 480), (1920, 1080), (1242, 375), (3840, 2160), (97, 131), (64, 128), (30, 30)])
def test_level_enumeration_equals_oracle(shape):
    port = Port()
    W, H = shape
    for win in [(64, 128), (48, 96)]:
        for scale0, nlevels in [(1.05, 64), (1.02, 128), (1.05, 15), (1.0, 64), (0.9, 5), (1.3, 3), (1.05, 1)]:
            a = HOG.enumerateLevels(W, H, win, scale0, nlevels)
            b = port.levels(W, H, win, scale0, nlevels)
            for x, y in zip(a, b):
                assert np.array_equal(x, y)
    # Appendix B of SURVEY.md
    assert len(HOG.enumerateLevels(1920, 1080)[0]) == 44
    assert len(HOG.enumerateLevels(640, 480)[0]) == 28
    assert len(HOG.enumerateLevels(3840, 2160)[0]) == 58
    assert len(HOG.enumerateLevels(1920, 1080, scale0=1.02, nlevels=128)[0]) == 108


def test_group_rectangles_equals_oracle():
    port = Port()
    rng = np.random.RandomState(1)
    for n, thr in [(0, 2), (1, 1), (1, 0), (50, 1), (300, 2), (2000, 2), (400, 0), (400, 5), (3000, 3)]:
        centers = rng.randint(0, 900, size=(max(n // 10, 1), 2))
        idx = rng.randint(0, len(centers), size=n)
        wh = rng.choice([64, 67, 71, 74, 78, 128, 135], size=n)
        r = np.stack([centers[idx, 0] + rng.randint(-12, 13, n), centers[idx, 1] + rng.randint(-12, 13, n),
                      wh, wh * 2], 1).astype(np.int32).reshape(-1, 4)
        a = HOG.groupRectangles(r, thr)
        b = port.group_rectangles(r, thr)
        assert np.array_equal(a, b), (n, thr)
        if Ref.available() and n:
            assert np.array_equal(a, Ref().group_rectangles(r, thr))


@pytest.mark.skipif(not Ref.available(), reason="oracle/_ref/libcvref.so not built")
def test_group_rectangles_lattice_fuzz_vs_reference():
    """Tie-heavy inputs (rects on a lattice, a handful of sizes, eps at and around the similarity boundary): classes,
    their order, the float means and the inside-a-stronger-class filter equal the unmodified cv::groupRectangles."""
    ref = Ref()
    for seed in range(200):
        rng = np.random.RandomState(seed)
        n = int(rng.randint(1, 120))
        s = rng.choice([32, 40, 48, 64, 65, 80], n)
        r = np.stack([rng.randint(0, 20, n) * rng.choice([4, 8]), rng.randint(0, 12, n) * rng.choice([4, 8]), s, 2 * s], 1).astype(np.int32)
        for thr in (1, 2, 3):
            for eps in (0.2, 0.5, 0.05):
                assert np.array_equal(HOG.groupRectangles(r, thr, eps), ref.group_rectangles(r, thr, eps)), (seed, thr, eps)


def test_headers_are_plain_c(tmp_path):
    """The boundary is a C ABI: both headers must compile as C99 (no C++ types, no torch types in the signatures)."""
    import subprocess
    src = tmp_path / "abi.c"
    src.write_text('#include "hogb200.h"\n#include "tbdb200.h"\nint main(void) { hogb200_geom g; tbdb200_args a; (void)g; (void)a; return 0; }\n')
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-fsyntax-only", "-I" + os.path.join(ROOT, "include"), str(src)])


def test_group_rectangles_meanshift_equals_reference():
    """hogb200_group_rectangles_meanshift (host) against the unmodified cv::groupRectangles_meanshift: identical rects in
    identical order, mode densities to 1e-9 relative -- clustered hits over several pyramid scales, isolated hits, ties,
    a threshold that drops some modes, and the empty list."""
    from _oracle import Ref
    from opencv_b200 import HOG
    if not Ref.available():
        pytest.skip("oracle/_ref not present")
    ref = Ref()
    for seed in range(12):
        rng = np.random.RandomState(seed)
        n = int(rng.randint(1, 120))
        k = rng.randint(0, 4, n)
        lvl = rng.randint(0, 12, n)
        sc = 1.05 ** lvl
        cx = np.array([200, 520, 800, 330])[k] + rng.randint(-12, 13, n); cy = np.array([220, 300, 150, 400])[k] + rng.randint(-20, 21, n)
        w = np.rint(64 * sc).astype(np.int32); h = np.rint(128 * sc).astype(np.int32)
        rects = np.stack([cx - w // 2, cy - h // 2, w, h], 1).astype(np.int32)
        wts = rng.uniform(0.05, 2.0, n)
        if seed % 3 == 0:
            rects[n // 2:] = rects[:n - n // 2]; wts[n // 2:] = wts[:n - n // 2]; sc[n // 2:] = sc[:n - n // 2]     # exact duplicates
        for thr in (0.0, 0.02, 1.0):
            er, ew = ref.group_rectangles_meanshift(rects, wts, sc, thr)
            gr, gw = HOG.groupRectanglesMeanshift(rects, wts, sc, thr)
            assert np.array_equal(gr, er), (seed, thr)
            assert np.allclose(gw, ew, rtol=1e-9, atol=0)
    gr, gw = HOG.groupRectanglesMeanshift(np.zeros((0, 4), np.int32), [], [])
    assert len(gr) == 0 and len(gw) == 0
