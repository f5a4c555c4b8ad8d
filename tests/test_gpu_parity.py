"""GPU parity tests proper: the CUDA path, called through the C ABI (libhogb200.so), against the oracle
(oracle/hog_oracle.c, pinned to the unmodified reference by tests/test_oracle_vs_ref.py) on the same seeded inputs.

Bars (BASELINE.json north_star):
  * pyramid levels: BIT-EXACT to cv::resize INTER_LINEAR
  * descriptors / block histograms: within 1e-4 (relative L2 per block and max-abs, values are in [0, 1])
  * SVM scores: within 1e-3 absolute
  * hit lists / grouped detection sets: identical, except for windows whose score lies within 1e-3 of the threshold
"""
import numpy as np
import pytest

from _oracle import Port, Ref, synthetic_frame, parts, detections_match

pytestmark = pytest.mark.gpu

DESC_TOL = 1e-4
SCORE_TOL = 1e-3


@pytest.fixture(scope="module")
def port():
    return Port()


@pytest.fixture(scope="module")
def hog():
    from opencv_b200 import HOG
    h = HOG.create()
    h.setSVMDetector(h.getDefaultPeopleDetector())
    return h


def rel_l2(a, b):
    return float(np.linalg.norm(a.astype(np.float64) - b) / max(np.linalg.norm(b.astype(np.float64)), 1e-12))


# ------------------------------------------------------------------ pyramid
@pytest.mark.parametrize("cn", [1, 4])
@pytest.mark.parametrize("shape", [(640, 480), (1242, 375), (321, 243)])
def test_pyramid_levels_bit_exact(hog, port, cn, shape):
    W, H = shape
    img = synthetic_frame(100 + cn, W, H, cn, blur=False)
    sc, lw, lh = hog.levels(W, H)
    sizes = sorted(set(zip(lw.tolist(), lh.tolist())) | {(W // 2, H // 2), (W // 3, H // 3), (W - 1, H - 1), (17, 9)})
    for dw, dh in sizes:
        got = hog.resize(img, dw, dh)
        exp = port.resize(img, dw, dh)
        assert np.array_equal(got, exp), (dw, dh, int(np.abs(got.astype(int) - exp).max()))


def test_pyramid_exact_half_and_upscale(hog, port):
    img = synthetic_frame(7, 480, 270, 4, blur=False)
    for dw, dh in [(240, 135), (480, 270), (960, 540), (500, 300)]:
        assert np.array_equal(hog.resize(img, dw, dh), port.resize(img, dw, dh)), (dw, dh)


def test_pyramid_1080p_levels_bit_exact(hog, port):
    """Full BASELINE size, a sample of the 44 levels (the oracle resizes 1080p in milliseconds)."""
    img = synthetic_frame(5, 1920, 1080, 4)
    sc, lw, lh = hog.levels(1920, 1080)
    assert len(sc) == 44
    for i in [1, 2, 7, 15, 30, 43]:
        assert np.array_equal(hog.resize(img, int(lw[i]), int(lh[i])), port.resize(img, int(lw[i]), int(lh[i])))


def test_pitched_device_input(hog, port):
    import torch
    img = synthetic_frame(8, 333, 222, 4, blur=False)
    big = torch.zeros((240, 400, 4), dtype=torch.uint8, device="cuda")
    big[5:227, 16:349] = torch.from_numpy(img).cuda()
    roi = big[5:227, 16:349]                     # a pitched ROI, like a GpuMat sub-matrix
    assert np.array_equal(hog.resize(roi, 200, 133), port.resize(img, 200, 133))
    a = hog.blockHistograms(roi)
    b = hog.blockHistograms(img)
    assert np.array_equal(a, b)


# ------------------------------------------------------------------ block histograms / descriptors
@pytest.mark.parametrize("cn", [1, 4])
@pytest.mark.parametrize("shape", [(64, 128), (640, 480), (333, 222), (251, 517), (16, 16), (24, 40)])
def test_block_histograms(hog, port, cn, shape):
    W, H = shape
    img = synthetic_frame(200 + cn + W, W, H, cn)
    got = hog.blockHistograms(img)
    exp = port.block_hists(img)
    assert got.shape == exp.shape
    assert np.abs(got - exp).max() <= DESC_TOL, float(np.abs(got - exp).max())
    err = np.linalg.norm(got - exp, axis=2) / np.maximum(np.linalg.norm(exp, axis=2), 1e-6)
    assert err.max() <= DESC_TOL, float(err.max())


def test_block_histograms_noise_and_flat(hog, port):
    """Unblurred noise (every pixel a strong edge) and a flat image (all gradients zero)."""
    img = synthetic_frame(3, 320, 240, 4, blur=False)
    assert np.abs(hog.blockHistograms(img) - port.block_hists(img)).max() <= DESC_TOL
    flat = np.full((96, 160), 77, np.uint8)
    got = hog.blockHistograms(flat)
    assert np.array_equal(got, port.block_hists(flat)) and not np.isnan(got).any()


def test_gamma_off_and_sigma(port):
    from opencv_b200 import HOG
    from _oracle import DEFAULT_GEOM
    h = HOG.create()
    img = synthetic_frame(4, 200, 160, 1)
    h.setGammaCorrection(False)
    assert np.abs(h.blockHistograms(img) - port.block_hists(img, dict(DEFAULT_GEOM, gamma=False))).max() <= DESC_TOL
    h.setGammaCorrection(True); h.setWinSigma(3.0); h.setL2HysThreshold(0.3)
    assert h.getWinSigma() == 3.0
    exp = port.block_hists(img, dict(DEFAULT_GEOM, win_sigma=3.0, l2hys=0.3))
    assert np.abs(h.blockHistograms(img) - exp).max() <= DESC_TOL


@pytest.mark.parametrize("cn", [1, 4])
def test_compute_descriptors_both_formats(hog, port, cn):
    """HOG::compute vs HOGDescriptor::compute; ROW_BY_ROW is a block permutation of COL_BY_COL
    (cudaobjdetect/test/test_objdetect.cpp:173-219)."""
    img = synthetic_frame(31 + cn, 64 * 3, 128 * 2, cn)
    for ws in [(64, 128), (8, 8), (16, 8)]:
        hog.setWinStride(ws)
        hog.setDescriptorFormat(hog.DESCR_FORMAT_COL_BY_COL)
        col = hog.compute(img)
        exp = port.compute(img, ws)
        assert col.shape == exp.shape == (parts(192, 64, ws[0]) * parts(256, 128, ws[1]), 3780)
        assert np.abs(col - exp).max() <= DESC_TOL
        hog.setDescriptorFormat(hog.DESCR_FORMAT_ROW_BY_ROW)
        row = hog.compute(img)
        perm = row.reshape(-1, 15, 7, 36).transpose(0, 2, 1, 3).reshape(-1, 3780)
        assert np.array_equal(perm, col)
    hog.setWinStride((8, 8)); hog.setDescriptorFormat(hog.DESCR_FORMAT_COL_BY_COL)


def test_descriptors_vs_unmodified_reference(hog):
    """Directly against cv::HOGDescriptor::compute of the unmodified reference (when oracle/_ref travelled)."""
    if not Ref.available():
        pytest.skip("oracle/_ref not present")
    ref = Ref()
    img = ref.rng_fill(0x77, 128, 256, 1)[:, :, 0]     # the Strict test's kind of input
    got = hog.compute(img)
    exp = ref.compute(img)
    assert np.abs(got - exp).max() <= DESC_TOL and rel_l2(got, exp) <= DESC_TOL


# ------------------------------------------------------------------ scores / detect
@pytest.mark.parametrize("cn", [1, 4])
@pytest.mark.parametrize("shape", [(64, 128), (320, 272), (641, 483), (71, 135)])
def test_window_scores(hog, port, cn, shape):
    W, H = shape
    img = synthetic_frame(300 + cn + W, W, H, cn, silhouettes=2)
    det = hog.getDefaultPeopleDetector()
    got = hog.windowScores(img)
    exp = port.scores(img, det)
    assert got.shape == exp.shape
    assert np.abs(got - exp).max() <= SCORE_TOL, float(np.abs(got - exp).max())


def test_window_scores_stride16(port):
    from opencv_b200 import HOG
    h = HOG.create(); det = h.getDefaultPeopleDetector(); h.setSVMDetector(det)
    h.setWinStride((16, 24))
    img = synthetic_frame(9, 400, 300, 1)
    got = h.windowScores(img)
    exp = port.scores(img, det, (16, 24))
    assert got.shape == exp.shape and np.abs(got - exp).max() <= SCORE_TOL


def test_daimler_48x96(port):
    from opencv_b200 import HOG
    from _oracle import DEFAULT_GEOM
    h = HOG.create(win_size=(48, 96)); det = h.getDefaultPeopleDetector()
    assert det.size == 1981 and h.getDescriptorSize() == 1980
    h.setSVMDetector(det)
    img = synthetic_frame(10, 300, 260, 1)
    g = dict(DEFAULT_GEOM, win=(48, 96))
    assert np.abs(h.windowScores(img) - port.scores(img, det, (8, 8), g)).max() <= SCORE_TOL
    assert np.abs(h.compute(img) - port.compute(img, (8, 8), g)).max() <= DESC_TOL


@pytest.mark.parametrize("win", [(128, 64), (32, 32), (96, 160), (16, 16), (176, 48), (72, 72), (64, 256)])
def test_any_window_size(port, win):
    """cv::cuda::HOG takes any block-aligned window (cudaobjdetect.hpp:88-92, ctor asserts hog.cpp:252-264, classify kernels
    hog.cu:394-514 loop over nblocks_win_x/y): a wide "vehicle" window (128x64 -> 15 x 7 blocks, scored in two column
    chunks), square and tiny ones, 21 blocks wide, 31 blocks high.  Scores, descriptors and detections against the oracle."""
    from opencv_b200 import HOG
    from _oracle import DEFAULT_GEOM, desc_size
    g = dict(DEFAULT_GEOM, win=win)
    n = desc_size(g)
    rng = np.random.RandomState(win[0] * 1000 + win[1])
    det = (rng.randn(n + 1) * (3.0 / np.sqrt(n))).astype(np.float32)
    det[n] = -0.25
    h = HOG.create(win_size=win)
    assert h.getDescriptorSize() == n
    h.setSVMDetector(det)
    for cn, shape in ((4, (413, 331)), (1, (300, 290))):
        img = synthetic_frame(win[0] + cn, shape[0], shape[1], cn, silhouettes=2)
        got = h.windowScores(img)
        exp = port.scores(img, det, (8, 8), g)
        assert got.shape == exp.shape and got.size > 0
        assert np.abs(got - exp).max() <= SCORE_TOL, float(np.abs(got - exp).max())
    small = synthetic_frame(3, win[0] + 24, win[1] + 16, 1)
    assert np.abs(h.compute(small) - port.compute(small, (8, 8), g)).max() <= DESC_TOL
    # the whole path with this window: levels end where the window no longer fits, rects scale with the window
    img = synthetic_frame(win[1], 500, 420, 4, silhouettes=3)
    thr = float(np.percentile(port.scores(img, det, (8, 8), g), 97))
    h.setHitThreshold(thr); h.setGroupThreshold(0)
    got = h.detectMultiScale(img)
    exp, _ = port.detect_multiscale(img, det, hit_threshold=thr, group_threshold=0, g=g)
    if not np.array_equal(got, exp):
        lo, _ = port.detect_multiscale(img, det, hit_threshold=thr - SCORE_TOL, group_threshold=0, g=g)
        hi, _ = port.detect_multiscale(img, det, hit_threshold=thr + SCORE_TOL, group_threshold=0, g=g)
        assert set(map(tuple, hi)) <= set(map(tuple, got)) <= set(map(tuple, lo))
    assert len(got) > 0


def test_two_wide_detectors_one_pass(port):
    """Two 128x64 detectors (pedestrian + vehicle in the north star's words, here both synthetic): one pass over the
    histogram ring scores both; each equals its own single-detector run and the oracle."""
    from opencv_b200 import HOG
    from _oracle import DEFAULT_GEOM
    g = dict(DEFAULT_GEOM, win=(128, 64))
    rng = np.random.RandomState(77)
    dets = [(rng.randn(15 * 7 * 36 + 1) * 0.05).astype(np.float32) for _ in range(3)]
    h = HOG.create(win_size=(128, 64))
    for i, d in enumerate(dets):
        h.setSVMDetector(d, slot=i)
    img = synthetic_frame(5, 640, 360, 4, silhouettes=3)
    for i, d in enumerate(dets):
        assert np.abs(h.windowScores(img, slot=i) - port.scores(img, d, (8, 8), g)).max() <= SCORE_TOL
    h.setHitThreshold(-0.5); h.setGroupThreshold(0)
    out = h.detectMultiScaleBatch([img, img], nslots=3)
    for i, d in enumerate(dets):
        ok, one, exp = detections_match(h, port, img, d, -0.5, 0, slot=i, g=g)
        assert ok and len(one) > 0 and np.array_equal(out[i], one) and np.array_equal(out[3 + i], one)


def test_oversized_window_is_refused():
    """Windows the scoring kernel cannot hold are refused by every scoring entry point before anything is queued
    (batch included), never mis-computed; compute() and blockHistograms() still work for them."""
    from opencv_b200 import HOG, HogError
    img = synthetic_frame(1, 640, 480, 1)
    for win in ((256, 256), (184, 64)):
        n = ((win[0] - 16) // 8 + 1) * ((win[1] - 16) // 8 + 1) * 36
        h = HOG.create(win_size=win)
        h.setSVMDetector(np.zeros(n + 1, np.float32))
        for call in (lambda: h.detect(img), lambda: h.detectMultiScale(img), lambda: h.detectMultiScaleBatch([img, img]),
                     lambda: h.windowScores(img)):
            with pytest.raises(HogError) as e:
                call()
            assert e.value.status == 4
        assert h.blockHistograms(img).shape == (59, 79, 36)
        h.setWinStride((64, 64))
        assert h.compute(img).shape[1] == n


def _hits_match(got_xy, exp_scores, thr, stride=8):
    exp_hit = exp_scores >= thr
    near = np.abs(exp_scores - thr) <= SCORE_TOL
    got_mask = np.zeros_like(exp_hit)
    got_mask[got_xy[:, 1] // stride, got_xy[:, 0] // stride] = True
    bad = (got_mask != exp_hit) & ~near
    return not bad.any()


def test_detect_hits_and_confidences(hog, port):
    img = synthetic_frame(12, 480, 360, 4, silhouettes=3)
    det = hog.getDefaultPeopleDetector()
    exp = port.scores(img, det)
    thr = float(np.percentile(exp, 90))
    hog.setHitThreshold(thr)
    hits = hog.detect(img)
    assert _hits_match(hits, exp, thr)
    # row-major order (hog.cpp:373-379)
    key = hits[:, 1].astype(np.int64) * 100000 + hits[:, 0]
    assert np.all(np.diff(key) > 0)
    hits2, conf = hog.detect(img, confidences=True)
    assert np.array_equal(hits, hits2)
    assert np.abs(conf - exp[hits2[:, 1] // 8, hits2[:, 0] // 8]).max() <= SCORE_TOL
    hog.setHitThreshold(0.0)


def test_empty_detector_and_small_image(port):
    from opencv_b200 import HOG, HogError
    h = HOG.create()
    img = synthetic_frame(1, 200, 200, 1)
    assert len(h.detect(img)) == 0 and len(h.detectMultiScale(img)) == 0          # hog.cpp:343-345
    h.setSVMDetector(h.getDefaultPeopleDetector())
    small = synthetic_frame(2, 40, 60, 1)                                           # smaller than the window
    assert len(h.detect(small)) == 0 and len(h.detectMultiScale(small)) == 0
    assert h.compute(small).shape == (0, 3780)
    with pytest.raises(HogError):
        h.detect(np.zeros((100, 100, 2), np.uint8))                                 # hog.cpp:340 (8UC3 is widened, see below)
    h.setWinStride((12, 8))
    with pytest.raises(HogError):
        h.detect(img)                                                               # hog.cpp:341
    h.setWinStride((8, 8)); h.setGroupThreshold(2)
    with pytest.raises(HogError):
        h.detectMultiScale(img, confidences=True)                                   # hog.cpp:422
    with pytest.raises(HogError):
        h.setSVMDetector(np.zeros(100, np.float32))                                 # hog.cpp:301
    with pytest.raises(HogError):
        HOG.create(win_size=(64, 128), block_size=(16, 16), block_stride=(4, 4))    # hog.cpp:263


# ------------------------------------------------------------------ detectMultiScale
def _set_equal_mod_threshold(hog, port, img, det, hit, group, scale0=1.05, nlevels=64, slot=0):
    """tests/_oracle.py detections_match: identical grouped sets, or near-threshold candidates only + regroup check."""
    return detections_match(hog, port, img, det, hit, group, scale0, nlevels, slot)


@pytest.mark.parametrize("cn,shape", [(1, (640, 480)), (4, (640, 480)), (4, (1242, 375)), (1, (300, 400))])
def test_detect_multiscale_candidates_and_groups(hog, port, cn, shape):
    W, H = shape
    img = synthetic_frame(400 + cn + W, W, H, cn, silhouettes=5)
    det = hog.getDefaultPeopleDetector()
    for hit, group in [(-1.5, 0), (-1.5, 2), (-2.5, 2), (0.0, 2)]:
        ok, got, exp = _set_equal_mod_threshold(hog, port, img, det, hit, group)
        assert ok, (hit, group, len(got), len(exp))
    hog.setHitThreshold(0.0); hog.setGroupThreshold(2)


def test_detect_multiscale_confidences_and_order(hog, port):
    img = synthetic_frame(77, 500, 400, 4, silhouettes=4)
    det = hog.getDefaultPeopleDetector()
    hog.setHitThreshold(-2.0); hog.setGroupThreshold(0)
    r, c = hog.detectMultiScale(img, confidences=True)
    er, ec = port.detect_multiscale(img, det, hit_threshold=-2.0, group_threshold=0)
    if len(r) == len(er):
        assert np.array_equal(r, er) and np.abs(c - ec).max() <= SCORE_TOL      # same order: level, row-major
    else:       # only windows within SCORE_TOL of the threshold may differ; the common ones keep order and confidence
        ok, _, _ = _set_equal_mod_threshold(hog, port, img, det, -2.0, 0)
        assert ok
        common = set(map(tuple, r)) & set(map(tuple, er))
        ri = [i for i, x in enumerate(map(tuple, r)) if x in common]; ei = [i for i, x in enumerate(map(tuple, er)) if x in common]
        assert np.array_equal(r[ri], er[ei]) and np.abs(c[ri] - ec[ei]).max() <= SCORE_TOL
    hog.setHitThreshold(0.0); hog.setGroupThreshold(2)


def test_sample_settings_tbd(hog, port):
    """example_gpu_tbd's settings: nlevels=15, hit_threshold=0.45, group 2 (samples/gpu/tbd.cpp:437-442)."""
    img = synthetic_frame(55, 1242, 375, 4, silhouettes=6)
    det = hog.getDefaultPeopleDetector()
    ok, got, exp = _set_equal_mod_threshold(hog, port, img, det, 0.45, 2, 1.05, 15)
    assert ok
    ok, got, exp = _set_equal_mod_threshold(hog, port, img, det, -1.0, 2, 1.05, 15)
    assert ok and len(exp) > 0
    hog.setHitThreshold(0.0); hog.setGroupThreshold(2); hog.setNumLevels(64)


def test_two_detectors_and_batch(hog, port):
    """Pedestrian + (synthetic) vehicle detector scored in one pass; batch API == per-frame API."""
    rng = np.random.RandomState(5)
    veh = (rng.randn(3781) * 0.05).astype(np.float32); veh[3780] = -0.5
    hog.setSVMDetector(veh, slot=1)
    det = hog.getDefaultPeopleDetector()
    hog.setHitThreshold(-1.0); hog.setGroupThreshold(2)
    frames = [synthetic_frame(600 + i, 1242, 375, 4, silhouettes=3) for i in range(5)]
    out = hog.detectMultiScaleBatch(frames, nslots=2)
    assert len(out) == 10
    for i, f in enumerate(frames):
        assert np.array_equal(out[2 * i], hog.detectMultiScale(f, slot=0))
        assert np.array_equal(out[2 * i + 1], hog.detectMultiScale(f, slot=1))
    for slot, d in ((0, det), (1, veh)):
        ok, got, _ = _set_equal_mod_threshold(hog, port, frames[0], d, -1.0, 2, slot=slot)
        assert ok and np.array_equal(got, out[slot])
    import torch
    dev = [torch.from_numpy(f).cuda() for f in frames]
    out2 = hog.detectMultiScaleBatch(dev, nslots=2)
    assert all(np.array_equal(a, b) for a, b in zip(out, out2))
    hog.setHitThreshold(0.0)


@pytest.mark.parametrize("shape", [(640, 480), (321, 243), (1242, 375)])
def test_bgr_input_equals_bgra(hog, port, shape):
    """8UC3 frames (what example_gpu_tbd reads before its host cvtColor, samples/gpu/tbd.cpp:573-599) are widened on
    the device: every result equals the one of the 8UC4 frame, host or device input, odd widths and pitches too."""
    import torch
    w, h = shape
    bgra = synthetic_frame(900 + w, w, h, 4, silhouettes=3)
    bgr = np.ascontiguousarray(bgra[:, :, :3])
    hog.setHitThreshold(-1.5); hog.setGroupThreshold(0)
    r4, c4 = hog.detectMultiScale(bgra, confidences=True)
    r3, c3 = hog.detectMultiScale(bgr, confidences=True)
    assert len(r4) > 0 and np.array_equal(r3, r4) and np.array_equal(c3, c4)
    r3d, c3d = hog.detectMultiScale(torch.from_numpy(bgr).cuda(), confidences=True)
    assert np.array_equal(r3d, r4) and np.array_equal(c3d, c4)
    wide = np.zeros((h, w + 5, 3), np.uint8); wide[:, :w] = bgr          # a pitched view (pitch = 3 * (w + 5))
    r3p, c3p = hog.detectMultiScale(torch.from_numpy(wide).cuda()[:, :w], confidences=True)
    assert np.array_equal(r3p, r4) and np.array_equal(c3p, c4)
    assert np.array_equal(hog.blockHistograms(bgr), hog.blockHistograms(bgra))
    hog.setGroupThreshold(2)
    out = hog.detectMultiScaleBatch([bgr, bgra, bgr])
    assert np.array_equal(out[0], out[1]) and np.array_equal(out[0], out[2])
    ok, got, _ = _set_equal_mod_threshold(hog, port, bgra, hog.getDefaultPeopleDetector(), -1.5, 2)
    assert ok and np.array_equal(got, out[0])
    hog.setHitThreshold(0.0)


@pytest.mark.parametrize("shape", [(1242, 375, 4), (644, 480, 1), (333, 251, 4), (642, 300, 1)])
def test_batch_frames_with_unaligned_rows(hog, shape):
    """Batch path, frames whose rows are no multiple of 16 bytes: contiguous host frames go up as one flat copy and are
    re-pitched by a kernel on the compute stream, device frames are re-pitched where they lie (hog_host.cu stage_frame;
    rows that are no multiple of 4 bytes keep the 2-D copy): always the single-frame path's result, which stages through
    a 2-D copy; submit / collect included."""
    import torch
    w, h, cn = shape
    frames = [synthetic_frame(6100 + 7 * i + w, w, h, cn, silhouettes=3) for i in range(5)]
    dev = [torch.from_numpy(f).cuda() for f in frames]
    hog.setHitThreshold(-1.5); hog.setGroupThreshold(0)
    try:
        want = [hog.detectMultiScale(f) for f in frames]
        assert sum(len(x) for x in want) > 0
        for src in (frames, dev, [frames[0], dev[1], frames[2], dev[3], frames[4]]):
            got = hog.detectMultiScaleBatch(src)
            assert all(np.array_equal(a, b) for a, b in zip(got, want))
        t = hog.submitBatch(frames)
        got = hog.collectBatch(t)
        assert all(np.array_equal(a, b) for a, b in zip(got, want))
    finally:
        hog.setHitThreshold(0.0); hog.setGroupThreshold(2)


def test_save_load_detector_file(hog, tmp_path):
    """cv::cuda::HOG::save / load (cudaobjdetect/src/hog.cpp:599-679) with the detector included: a handle created from
    the file detects exactly what the original does; the reference's own writer output loads as well."""
    from opencv_b200 import HOG
    from _oracle import Ref
    img = synthetic_frame(31, 640, 480, 4, silhouettes=3)
    hog.setHitThreshold(-1.0); hog.setGroupThreshold(0); hog.setNumLevels(20)
    want = hog.detectMultiScale(img, confidences=True)
    for ext in ("yml", "xml"):
        path = str(tmp_path / ("people." + ext))
        hog.save(path)
        h2 = HOG.load(path)
        assert h2.getNumLevels() == 20 and h2.getWinSigma() == 4.0 and h2.getGammaCorrection() is True
        h2.setHitThreshold(-1.0); h2.setGroupThreshold(0)
        got = h2.detectMultiScale(img, confidences=True)
        assert len(want[0]) > 0 and np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    if Ref.available():
        ref = Ref()
        path = str(tmp_path / "from_reference.yml")
        ref.hog_save(path, ref.people_detector(0), nlevels=20)
        h3 = HOG.load(path)
        h3.setHitThreshold(-1.0); h3.setGroupThreshold(0)
        got = h3.detectMultiScale(img, confidences=True)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    hog.setHitThreshold(0.0); hog.setGroupThreshold(2); hog.setNumLevels(64)


def test_dense_candidates_overflow_path(hog, port):
    """hit_threshold far below every score: every window is a candidate (dense-pyramid stress, config 5 in
    small): exercises the candidate-list regrow path and the ordering."""
    img = synthetic_frame(88, 640, 480, 1)
    hog.setHitThreshold(-100.0); hog.setGroupThreshold(0)
    r = hog.detectMultiScale(img)
    sc, lw, lh = hog.levels(640, 480)
    total = sum(parts(int(w), 64, 8) * parts(int(h), 128, 8) for w, h in zip(lw, lh))
    assert len(r) == total == 25132
    er, _ = port.detect_multiscale(img, hog.getDefaultPeopleDetector(), hit_threshold=-100.0, group_threshold=0)
    assert np.array_equal(r, er)
    hog.setHitThreshold(0.0); hog.setGroupThreshold(2)


def test_determinism(hog):
    img = synthetic_frame(99, 800, 600, 4, silhouettes=3)
    a = hog.blockHistograms(img); b = hog.blockHistograms(img)
    assert np.array_equal(a, b)
    assert np.array_equal(hog.windowScores(img), hog.windowScores(img))


def test_full_size_1080p_properties(hog, port):
    """BASELINE config 2 at full size through size-independent properties + a spot check against the oracle."""
    img = synthetic_frame(2024, 1920, 1080, 4, silhouettes=8)
    hog.setHitThreshold(-100.0); hog.setGroupThreshold(0)
    r, c = hog.detectMultiScale(img, confidences=True)
    assert len(r) == 261558                                    # every window of the 44 levels (Appendix B)
    sc, lw, lh = hog.levels(1920, 1080)
    # the level-0 part of the dense list must equal windowScores, in row-major order
    n0 = parts(1920, 64, 8) * parts(1080, 128, 8)
    ws = hog.windowScores(img)
    assert np.array_equal(c[:n0].astype(np.float32).reshape(ws.shape), ws)
    # whole levels against the oracle end to end (resize -> hist -> score): the base frame, a mid-pyramid level, the last one
    for lvl in (0, 20, 43):
        limg = port.resize(img, int(lw[lvl]), int(lh[lvl])) if lvl else img
        exp = port.scores(limg, hog.getDefaultPeopleDetector())
        off = sum(parts(int(w), 64, 8) * parts(int(h), 128, 8) for w, h in zip(lw[:lvl], lh[:lvl]))
        got = c[off:off + exp.size].reshape(exp.shape)
        assert np.abs(got - exp).max() <= SCORE_TOL
    hog.setHitThreshold(0.0); hog.setGroupThreshold(2)


def _dense_full_size(hog, port, img, scale, nlevels, expect_windows, spot_levels):
    """Every window of every level (hit threshold below all scores, no grouping): the count, the level-0 scores
    against windowScores, and whole levels end to end (resize -> histograms -> scores) against the oracle."""
    H, W = img.shape[:2]
    det = hog.getDefaultPeopleDetector()
    hog.setScaleFactor(scale); hog.setNumLevels(nlevels); hog.setHitThreshold(-100.0); hog.setGroupThreshold(0)
    try:
        r, c = hog.detectMultiScale(img, confidences=True)
        sc, lw, lh = hog.levels(W, H)
        per_level = [parts(int(w), 64, 8) * parts(int(h), 128, 8) for w, h in zip(lw, lh)]
        assert len(r) == sum(per_level) == expect_windows
        ws = hog.windowScores(img)
        assert np.array_equal(c[:per_level[0]].astype(np.float32).reshape(ws.shape), ws)
        # rects of a level: cvRound(x*s), cvRound(y*s), cvRound(64*s), cvRound(128*s) in row-major window order
        for lvl in spot_levels:
            off = sum(per_level[:lvl])
            limg = port.resize(img, int(lw[lvl]), int(lh[lvl])) if lvl else img
            exp = port.scores(limg, det)
            got = c[off:off + exp.size].reshape(exp.shape)
            assert np.abs(got - exp).max() <= SCORE_TOL
            s = float(sc[lvl])
            assert tuple(r[off]) == (0, 0, int(np.rint(64 * s)), int(np.rint(128 * s)))
            assert tuple(r[off + 1][:2]) == (int(np.rint(8 * s)), 0)
    finally:
        hog.setScaleFactor(1.05); hog.setNumLevels(64); hog.setHitThreshold(0.0); hog.setGroupThreshold(2)


def test_full_size_4k_properties(hog, port):
    """BASELINE config 4 (3840x2160 8UC4): 58 levels, 1 209 184 windows."""
    img = synthetic_frame(4000, 3840, 2160, 4, silhouettes=8)
    _dense_full_size(hog, port, img, 1.05, 64, 1209184, [0, 31, 57])


def test_full_size_dense_pyramid_properties(hog, port):
    """BASELINE config 5 (1080p, scale 1.02, nlevels 128, hit 0 -> all-pass, group 0): 108 levels, 623 442
    candidate windows through the compaction / regrow / ordering path."""
    img = synthetic_frame(5000, 1920, 1080, 4, silhouettes=8)
    _dense_full_size(hog, port, img, 1.02, 128, 623442, [0, 53, 107])


def test_full_size_kitti_sequence_two_detectors(hog, port):
    """BASELINE config 3 in the sample's shape: a 1242x375 sequence, pedestrian + (synthetic) vehicle detector, the
    sample's nlevels=15 and nlevels=64: the batch path over a ragged group count equals the per-frame path, and the
    per-frame path equals the oracle on the first frame."""
    rng = np.random.RandomState(11)
    veh = (rng.randn(3781) * 0.05).astype(np.float32); veh[3780] = -0.5
    det = hog.getDefaultPeopleDetector()
    hog.setSVMDetector(veh, slot=1)
    frames = [synthetic_frame(7000 + i, 1242, 375, 4, silhouettes=4) for i in range(11)]    # 2 groups of 4 + 3
    try:
        for nlevels, nlv in ((15, 15), (64, 23)):
            hog.setNumLevels(nlevels); hog.setHitThreshold(-0.75); hog.setGroupThreshold(2)
            assert len(hog.levels(1242, 375)[0]) == nlv
            out = hog.detectMultiScaleBatch(frames, nslots=2)
            assert len(out) == 22
            for i, f in enumerate(frames):
                assert np.array_equal(out[2 * i], hog.detectMultiScale(f, slot=0))
                assert np.array_equal(out[2 * i + 1], hog.detectMultiScale(f, slot=1))
            for slot, d in ((0, det), (1, veh)):
                ok, got, exp = _set_equal_mod_threshold(hog, port, frames[0], d, -0.75, 2, 1.05, nlevels, slot=slot)
                assert ok
    finally:
        hog.setNumLevels(64); hog.setHitThreshold(0.0); hog.setGroupThreshold(2)


def test_batch_tail_on_device_equals_host_tail(hog, port, monkeypatch):
    """The batch entry point's tail (reference order, rect scaling, groupRectangles per (frame, detector)) computed on
    the device for the whole chunk (group.cu cand_group_device) against the host tail, forced either way: identical
    rects in identical order, grouped and ungrouped, one and two detectors, ragged groups; and against the oracle."""
    rng = np.random.RandomState(21)
    veh = (rng.randn(3781) * 0.05).astype(np.float32); veh[3780] = -0.5
    det = hog.getDefaultPeopleDetector()
    hog.setSVMDetector(veh, slot=1)
    frames = [synthetic_frame(9100 + i, 640, 480, 4, silhouettes=4) for i in range(11)]
    try:
        for hit, group, nslots in ((-1.5, 2, 1), (-1.5, 0, 2), (-2.5, 3, 2), (-0.5, 2, 2), (5.0, 2, 1)):
            hog.setHitThreshold(hit); hog.setGroupThreshold(group)
            monkeypatch.setenv("HOGB200_GPU_POST_MIN", "0")
            host = hog.detectMultiScaleBatch(frames, nslots=nslots)
            monkeypatch.setenv("HOGB200_GPU_POST_MIN", "1")
            dev = hog.detectMultiScaleBatch(frames, nslots=nslots)
            assert len(host) == len(dev) == len(frames) * nslots
            assert all(np.array_equal(a, b) for a, b in zip(host, dev)), (hit, group, nslots)
            assert hit > 0 or sum(len(a) for a in dev) > 0
            assert hog.lastTimings()["candidates"] > 0 or hit > 0
        monkeypatch.setenv("HOGB200_GPU_POST_MIN", "1")
        for slot, d in ((0, det), (1, veh)):
            ok, got, exp = detections_match(hog, port, frames[3], d, -1.5, 2, slot=slot)
            hog.setHitThreshold(-1.5); hog.setGroupThreshold(2)
            assert ok and np.array_equal(hog.detectMultiScaleBatch(frames[:5], nslots=2)[2 * 3 + slot], got)
    finally:
        hog.setHitThreshold(0.0); hog.setGroupThreshold(2)


def test_submit_collect_equals_blocking_batch(hog, monkeypatch):
    """hogb200_submit_batch / hogb200_collect_batch: two batches in flight, collected in and out of order, host and
    device frames, host and device tails, a candidate region that overflows (re-run by collect), a too small output
    array (results kept for the next collect): always the blocking entry point's output."""
    import torch
    from opencv_b200 import HogError
    a = [synthetic_frame(9300 + i, 640, 480, 4, silhouettes=4) for i in range(9)]
    b = [synthetic_frame(9400 + i, 1242, 375, 4, silhouettes=3) for i in range(6)]
    bd = [torch.from_numpy(f).cuda() for f in b]
    try:
        for hit, group, post in ((-1.5, 2, "0"), (-1.5, 2, "1"), (-2.0, 0, "1"), (3.0, 2, "1")):
            monkeypatch.setenv("HOGB200_GPU_POST_MIN", post)
            hog.setHitThreshold(hit); hog.setGroupThreshold(group)
            wa = hog.detectMultiScaleBatch(a); wb = hog.detectMultiScaleBatch(b)
            t0 = hog.submitBatch(a); t1 = hog.submitBatch(bd)
            with pytest.raises(HogError):
                hog.submitBatch(a)                                   # a third batch has to wait
            gb = hog.collectBatch(t1); ga = hog.collectBatch(t0)     # out of order
            assert all(np.array_equal(x, y) for x, y in zip(ga, wa)) and all(np.array_equal(x, y) for x, y in zip(gb, wb))
            t0 = hog.submitBatch(a + b)                              # two shapes in one batch: two chunks
            t1 = hog.submitBatch(b)
            g0 = hog.collectBatch(t0, capacity=4); g1 = hog.collectBatch(t1)
            assert all(np.array_equal(x, y) for x, y in zip(g0, wa + wb)) and all(np.array_equal(x, y) for x, y in zip(g1, wb))
        # every window a candidate: 9 VGA frames x 25 132 windows < the 262 144-candidate region; 1242x375 x 6 = 201 216 too --
        # 4K-sized demand comes from 64 frames: here force the overflow with a 40-frame VGA batch (1 005 280 candidates)
        hog.setHitThreshold(-100.0); hog.setGroupThreshold(0)
        many = [a[i % len(a)] for i in range(40)]
        want = hog.detectMultiScaleBatch(many, capacity=1 << 21)
        got = hog.collectBatch(hog.submitBatch(many), capacity=1 << 21)
        assert sum(len(x) for x in got) == 40 * 25132 and all(np.array_equal(x, y) for x, y in zip(got, want))
    finally:
        hog.setHitThreshold(0.0); hog.setGroupThreshold(2)


def test_async_and_roi_edge_cases(hog, port):
    """Edge cases of the entry points added in round 2: empty batches, bad tickets, 8UC3 frames through the ROI path,
    a handle without detector, parameters captured at submit time."""
    from opencv_b200 import HOG, HogError
    t = hog.submitBatch([])
    assert hog.collectBatch(t) == []
    with pytest.raises(HogError):
        hog.collectBatch(t)                                          # already collected
    with pytest.raises(HogError):
        hog._ck(hog._L.hogb200_collect_batch(hog._h, 7, None, 0, None))
    bgra = synthetic_frame(77, 320, 256, 4, silhouettes=2)
    bgr = np.ascontiguousarray(bgra[:, :, :3])
    loc = np.stack([np.arange(-1, 31) * 8, (np.arange(-1, 31) * 5 % 17) * 8], 1).astype(np.int32)
    loc = np.concatenate([loc, np.array([[400, 0], [0, 144], [-16, 8]], np.int32)])       # three outside the padded image
    hog.setHitThreshold(-3.0)
    f4, c4 = hog.detectROI(bgra, loc)
    f3, c3 = hog.detectROI(bgr, loc)
    assert np.array_equal(f3, f4) and np.array_equal(c3, c4) and (c4 == -10.0).any() and (c4 != -10.0).any()
    r4, _ = hog.detectMultiScaleROI(bgra, [(1.0, loc), (1.21, loc), (3.0, loc)])   # the last level is smaller than the window
    r3, cc = hog.detectMultiScaleROI(bgr, [(1.0, loc), (1.21, loc), (3.0, loc)])
    assert np.array_equal(r3, r4) and (cc[2] == -10.0).all()
    empty = HOG.create()
    assert len(empty.detectROI(bgra, loc)[0]) == 0 and len(empty.detectMultiScaleROI(bgra, [(1.0, loc)])[0]) == 0
    # the tail of an asynchronous batch uses the group threshold it was submitted with
    frames = [synthetic_frame(9500 + i, 640, 480, 4, silhouettes=4) for i in range(3)]
    hog.setHitThreshold(-1.5); hog.setGroupThreshold(2)
    want = hog.detectMultiScaleBatch(frames)
    t = hog.submitBatch(frames)
    hog.setGroupThreshold(0)
    got = hog.collectBatch(t)
    assert all(np.array_equal(a, b) for a, b in zip(got, want)) and hog.getGroupThreshold() == 0
    hog.setHitThreshold(0.0); hog.setGroupThreshold(2)


# ------------------------------------------------------------------ ROI variants (row f4)
def _roi_locations(rng, w, h, n, pad=8, win=(64, 128)):
    """window origins on the 8-pixel lattice: inside the image, reaching into the padding, and a few outside it"""
    xs = rng.randint(-pad // 8 - 1, (w + pad - win[0]) // 8 + 3, n) * 8
    ys = rng.randint(-pad // 8 - 1, (h + pad - win[1]) // 8 + 3, n) * 8
    return np.stack([xs, ys], 1).astype(np.int32)


@pytest.mark.parametrize("cn,shape", [(1, (320, 240)), (4, (333, 251)), (4, (640, 480))])
def test_detect_roi_vs_unmodified_reference(hog, cn, shape):
    """cv::HOGDescriptor::detectROI of the unmodified reference (objdetect/src/hog.cpp:3549-3660), padding 8 x 8 and 0 x 0 and
    16 x 24: confidences within 1e-3 for every location (windows reaching into the reflect-101 padding included), -10 for
    the locations outside, the same hits modulo the threshold tolerance."""
    if not Ref.available():
        pytest.skip("oracle/_ref not present")
    ref = Ref()
    W, H = shape
    img = synthetic_frame(70 + cn + W, W, H, cn, silhouettes=3)
    det = hog.getDefaultPeopleDetector()
    rng = np.random.RandomState(W)
    for pad in ((8, 8), (0, 0), (16, 24)):
        loc = _roi_locations(rng, W, H, 300, max(pad))
        _, ec0 = ref.detect_roi(img, det, loc, hit_threshold=0.0, padding=pad)
        thr = float(np.percentile(ec0[ec0 != -10.0], 80))
        ef, ec = ref.detect_roi(img, det, loc, hit_threshold=thr, padding=pad)
        hog.setHitThreshold(thr)
        gf, gc = hog.detectROI(img, loc, padding=pad)
        assert np.array_equal(gc == -10.0, ec == -10.0) and (ec == -10.0).any() and (ec != -10.0).sum() > 100
        assert np.abs(gc - ec).max() <= SCORE_TOL, float(np.abs(gc - ec).max())
        near = np.abs(ec - thr) <= SCORE_TOL
        sure = set(map(tuple, loc[(ec >= thr) & ~near & (ec != -10.0)])); maybe = set(map(tuple, loc[near]))
        assert sure <= set(map(tuple, gf)) <= sure | maybe and len(gf) > 0 and len(ef) > 0
    hog.setHitThreshold(0.0)
    from opencv_b200 import HogError
    with pytest.raises(HogError):
        hog.detectROI(img, [[12, 8]])                       # not on the cache-stride lattice: the reference asserts too
    assert len(hog.detectROI(img, np.zeros((0, 2), np.int32))[0]) == 0


def test_detect_multiscale_roi(hog, port):
    """cv::HOGDescriptor::detectMultiScaleROI.  (a) scale 1 only -- no resize involved -- against the unmodified function
    end to end, grouped and ungrouped; (b) several scales against the reference's own detectROI applied to the mandated
    INTER_LINEAR levels (the oracle convention of DESIGN.md section 2), rects scaled as hog.cpp:3527-3532."""
    if not Ref.available():
        pytest.skip("oracle/_ref not present")
    ref = Ref()
    W, H = 640, 480
    img = synthetic_frame(4321, W, H, 4, silhouettes=5)
    det = hog.getDefaultPeopleDetector()
    rng = np.random.RandomState(9)
    loc1 = _roi_locations(rng, W, H, 600)
    _, ec0 = ref.detect_roi(img, det, loc1, hit_threshold=0.0, padding=(8, 8))
    hit = float(np.percentile(ec0[ec0 != -10.0], 70)) + 0.0123
    hog.setHitThreshold(hit)
    for group in (0, 1):
        er, ec = ref.detect_multiscale_roi(img, det, [(1.0, loc1)], hit_threshold=hit, group_threshold=group)
        gr, gc = hog.detectMultiScaleROI(img, [(1.0, loc1)], group_threshold=group)
        assert np.abs(gc[0] - ec[0]).max() <= SCORE_TOL
        if np.abs(ec[0] - hit).min() > SCORE_TOL:            # no window within tolerance of the threshold: identical sets
            assert len(er) > 0 and sorted(map(tuple, gr)) == sorted(map(tuple, er))
    scales = [1.0, 1.05 ** 3, 1.5, 2.0, 2.6]
    rois, exp_rects = [], []
    for s in scales:
        lw, lh = int(np.rint(W / s)), int(np.rint(H / s))
        loc = _roi_locations(rng, lw, lh, 200)
        rois.append((s, loc))
        limg = port.resize(img, lw, lh) if (lw, lh) != (W, H) else img
        ef, ec = ref.detect_roi(limg, det, loc, hit_threshold=hit, padding=(8, 8))
        exp_rects.append((ef, ec, s))
    gr, gc = hog.detectMultiScaleROI(img, rois, group_threshold=0)
    sure, maybe = [], set()
    for (ef, ec, s), c, (_, loc) in zip(exp_rects, gc, rois):
        assert np.abs(c - ec).max() <= SCORE_TOL
        rect = lambda x, y: (int(np.rint(x * s)), int(np.rint(y * s)), int(np.rint(64 * s)), int(np.rint(128 * s)))
        for (x, y), e in zip(loc, ec):
            if e == -10.0:
                continue
            if abs(e - hit) <= SCORE_TOL:
                maybe.add(rect(x, y))                       # within tolerance of the threshold: either answer is right
            elif e >= hit:
                sure.append(rect(x, y))
    got = [tuple(r) for r in gr]
    assert len(sure) > 0 and [r for r in got if r not in maybe] == [r for r in sure if r not in maybe]   # roi order, location order
    assert set(got) <= set(sure) | maybe
    gg, _ = hog.detectMultiScaleROI(img, rois, group_threshold=1)
    assert np.array_equal(gg, port.group_rectangles(gr, 1))
    hog.setHitThreshold(0.0)


# ------------------------------------------------------------------ groupRectangles on the device (row f4)
def _rect_cloud(seed, n, nclusters, W=1920, H=1080):
    """Window-like rects (1:2) scattered around cluster centres at several scales, plus isolated ones, in random order."""
    rng = np.random.RandomState(seed)
    cx = rng.randint(0, W, nclusters); cy = rng.randint(0, H, nclusters); cw = rng.randint(40, 400, nclusters)
    k = rng.randint(0, nclusters, n)
    w = np.maximum(8, (cw[k] * rng.uniform(0.8, 1.25, n)).astype(np.int32))
    jit = np.maximum(1, cw[k] // 6)
    x = cx[k] + rng.randint(-1, 2, n) * rng.randint(0, 1 << 30, n) % jit
    y = cy[k] + rng.randint(-1, 2, n) * rng.randint(0, 1 << 30, n) % jit
    r = np.stack([x, y, w, 2 * w], 1).astype(np.int32)
    lone = rng.rand(n) < 0.05
    r[lone, 0] = rng.randint(0, W, lone.sum()); r[lone, 1] = rng.randint(0, H, lone.sum())
    return r


@pytest.mark.parametrize("n,nclusters", [(1, 1), (40, 3), (3000, 40), (50000, 300)])
def test_group_rectangles_device_equals_host(hog, n, nclusters):
    """Bit-identical rects in identical order: device union-find grouping vs the host restatement of
    cv::groupRectangles (itself pinned to the unmodified reference in tests/test_oracle_vs_ref.py / test_cabi_cpu.py)."""
    from opencv_b200 import HOG
    r = _rect_cloud(n, n, nclusters)
    for thr in (1, 2, 5):
        for eps in (0.2, 0.5):
            want = HOG.groupRectangles(r, thr, eps)
            got = hog.groupRectanglesDevice(r, thr, eps)
            assert np.array_equal(got, want), (n, thr, eps, len(got), len(want))
    assert len(hog.groupRectanglesDevice(np.zeros((0, 4), np.int32), 2)) == 0
    assert np.array_equal(hog.groupRectanglesDevice(r, 0), r)          # threshold <= 0: untouched (cascadedetect.cpp:66-75)


def test_group_rectangles_device_vs_reference():
    """Against the unmodified cv::groupRectangles directly."""
    from opencv_b200 import HOG
    h = HOG.create()
    ref = Ref()
    r = _rect_cloud(5, 20000, 150)
    assert np.array_equal(h.groupRectanglesDevice(r, 2, 0.2), ref.group_rectangles(r, 2, 0.2))


def test_dense_hits_grouped_on_device(hog, port):
    """A low hit threshold with grouping on: thousands of candidates per frame take the device grouping path inside
    detectMultiScale (>= 1024 candidates) and must give the oracle's grouped set."""
    img = synthetic_frame(321, 800, 600, 4, silhouettes=6)
    det = hog.getDefaultPeopleDetector()
    hog.setHitThreshold(-3.0); hog.setGroupThreshold(0)
    assert len(hog.detectMultiScale(img)) >= 2048
    ok, got, exp = _set_equal_mod_threshold(hog, port, img, det, -3.0, 2)
    assert ok and len(exp) > 0
    hog.setHitThreshold(0.0); hog.setGroupThreshold(2)


# ------------------------------------------------------------------ detector -> tracker (example_gpu_tbd's loop)
def _moving_scene(nframes, W=640, H=480):
    """A fixed textured background with three bright vertical ellipses walking across it."""
    base = synthetic_frame(4242, W, H, 4).astype(np.float32)
    yy, xx = np.ogrid[:H, :W]
    walkers = [(80, 150, 7, 1, 150), (520, 300, -9, -2, 110), (300, 240, 4, 0, 190)]     # x, y, vx, vy, height
    frames = []
    for k in range(nframes):
        img = base.copy()
        for (x, y, vx, vy, hh) in walkers:
            cx, cy, ww = x + vx * k, y + vy * k, hh / 3.0
            m = ((yy - cy) / (hh / 2.0)) ** 2 + ((xx - cx) / (ww / 2.0)) ** 2 <= 1
            img[m, :3] = img[m, :3] * 0.2 + 200
        frames.append(np.clip(np.rint(img), 0, 255).astype(np.uint8))
    return frames


def test_detector_feeds_tracker_like_the_sample(hog, port):
    """samples/gpu/tbd.cpp:596-700 end to end: detectMultiScale (sample settings: nlevels 15, group 2) -> one
    Detection(-1, frame, rect, -1.0) per grouped rect -> Tracker::performTrackingStep.  This pipeline (GPU detector +
    tbdb200 tracker) against the oracle detector + the UNMODIFIED reference tracker."""
    from opencv_b200 import tbd
    from test_tbd import RefTracker, our_state, same_state
    det = hog.getDefaultPeopleDetector()
    args = tbd.TbdArgs(10.0, 16, 4, 0.3, -1.0, 1)
    ours, ref = tbd.Tracker(args), RefTracker(args)
    ours.setBounds(0, 640, 0, 480)       # the reference hard-codes 1280x720; these frames are smaller than that
    hit = -1.6
    hog.setNumLevels(15); hog.setHitThreshold(hit); hog.setGroupThreshold(2)
    try:
        frames = _moving_scene(10)
        out = hog.detectMultiScaleBatch(frames)
        seen = 0
        for k, (img, rects) in enumerate(zip(frames, out)):
            ok, got, exp = _set_equal_mod_threshold(hog, port, img, det, hit, 2, 1.05, 15)
            assert ok and np.array_equal(got, rects)                   # batch path == per-frame path == oracle
            seen += len(rects)
            d_ours = [tbd.Detection(-1, k, tuple(int(v) for v in r), -1.0) for r in rects]
            d_ref = [tbd.Detection(-1, k, tuple(int(v) for v in r), -1.0) for r in (exp if len(exp) == len(got) else got)]
            ours.performTrackingStep(d_ours, k)
            ref.step(d_ref, k)
            if np.array_equal(got, exp):
                assert same_state(our_state(ours), ref.state()), k
        assert seen > 0
    finally:
        hog.setNumLevels(64); hog.setHitThreshold(0.0); hog.setGroupThreshold(2)


def test_api_edge_cases_of_the_reference(port, tmp_path):
    """behaviours of cv::cuda::HOG a drop-in has to keep."""
    from opencv_b200 import HOG, HogError
    h = HOG.create()
    det = h.getDefaultPeopleDetector()
    h.setSVMDetector(det)
    assert h.getWinSigma() == 4.0                                   # (block_w + block_h) / 8 while unset (hog.cpp:288-291)
    assert h.getDescriptorSize() == 3780 and h.getBlockHistogramSize() == 36
    assert (h.getNumLevels(), h.getHitThreshold(), h.getScaleFactor(), h.getGroupThreshold()) == (64, 0.0, 1.05, 2)
    # scale0 <= 1: exactly one level (hog.cpp:433) -- detectMultiScale degenerates to detect with 64x128 rects
    img = synthetic_frame(808, 320, 256, 4, silhouettes=3)
    h.setScaleFactor(1.0); h.setHitThreshold(-2.0); h.setGroupThreshold(0)
    assert len(h.levels(320, 256)[0]) == 1
    r = h.detectMultiScale(img)
    pts = h.detect(img)
    assert len(r) == len(pts) > 0 and np.array_equal(r[:, :2], pts) and (r[:, 2:] == (64, 128)).all()
    ok, got, _ = _set_equal_mod_threshold(h, port, img, det, -2.0, 0, scale0=1.0)
    assert ok and np.array_equal(got, r)
    # geometries outside block 16x16 / cell 8x8 / 9 bins are refused, not mis-computed (cudaobjdetect.hpp:89-93)
    for kw in (dict(win_size=(48, 96), block_size=(16, 16), block_stride=(8, 8), cell_size=(4, 4)),
               dict(win_size=(64, 128), block_size=(32, 32), block_stride=(16, 16), cell_size=(16, 16)),
               dict(nbins=8)):
        with pytest.raises(HogError) as e:
            HOG.create(**kw)
        assert e.value.status in (1, 4)
    # load() takes geometry, sigma, threshold, gamma, nlevels and the detector from the file and resets the run-time
    # parameters to their defaults (hog.cpp:599-630)
    h.setWinSigma(3.5); h.setNumLevels(17); h.setHitThreshold(1.25); h.setScaleFactor(1.3); h.setGroupThreshold(5)
    f = str(tmp_path / "det.yml")
    h.save(f)
    g = HOG.load(f)
    assert (g.getWinSigma(), g.getNumLevels()) == (3.5, 17)
    assert (g.getHitThreshold(), g.getScaleFactor(), g.getGroupThreshold(), g.getWinStride()) == (0.0, 1.05, 2, (8, 8))
    g.setHitThreshold(-2.0); g.setGroupThreshold(0); g.setWinSigma(-1.0)
    h.setWinSigma(-1.0)
    assert np.array_equal(g.detect(img), pts)                       # the detector travelled with the file


def test_group_rectangles_device_lattice_fuzz(hog):
    """Tie-heavy inputs (rects on a lattice, a handful of sizes, eps at and around the similarity boundary) through the
    device path: equal to the host restatement, which tests/test_cabi_cpu.py pins to the unmodified reference on the
    same generator."""
    from opencv_b200 import HOG
    for seed in range(60):
        rng = np.random.RandomState(seed)
        n = int(rng.randint(1, 400))
        s = rng.choice([32, 40, 48, 64, 65, 80], n)
        r = np.stack([rng.randint(0, 20, n) * rng.choice([4, 8]), rng.randint(0, 12, n) * rng.choice([4, 8]), s, 2 * s], 1).astype(np.int32)
        for thr in (1, 2, 3):
            for eps in (0.2, 0.5, 0.05):
                assert np.array_equal(hog.groupRectanglesDevice(r, thr, eps), HOG.groupRectangles(r, thr, eps)), (seed, thr, eps)
