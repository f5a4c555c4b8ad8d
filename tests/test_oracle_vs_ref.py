"""Pins the C oracle (oracle/hog_oracle.c) against the UNMODIFIED reference (oracle/_ref/libcvref.so).

CPU-only.  Skipped when oracle/_ref was not built (it is built in the build container from
/root/reference by oracle/Makefile and travels to the GPU box as a built .so).

The reference's own data-free pin for this path is Objdetect_HOGDetector_Strict
(modules/objdetect/test/test_cascadeandhog.cpp:1318-1346): random images rng.fill(UNIFORM,0,256) of size
k*winSize, 8UC1 and 8UC3, SIMD path == scalar re-implementation within FLT_EPSILON*100.  The same inputs are
used here (cv::RNG via the reference itself), and the oracle must be EQUAL to the reference's output, not
merely within that tolerance.
"""
import numpy as np
import pytest
from _oracle import Port, Ref, DEFAULT_GEOM, synthetic_frame, parts

pytestmark = pytest.mark.skipif(not Ref.available(), reason="oracle/_ref/libcvref.so not built")


@pytest.fixture(scope="module")
def libs():
    return Port(), Ref()


def test_reference_is_3_4_7(libs):
    _, ref = libs
    assert ref.version() == "3.4.7"
    det = ref.people_detector(0)
    assert det.size == 3781 and abs(det[3780] - (-6.66579151)) < 1e-6
    assert ref.people_detector(1).size == 1981


@pytest.mark.parametrize("shape", [(640, 480), (1920, 1080), (1242, 375), (97, 131)])
def test_levels_match_reference_sizes(libs, shape):
    # level sizes are checked indirectly: the reference resizes to exactly these sizes inside detectMultiScale;
    # here the rule itself (hog.cpp:424-441) is restated in numpy with Python's round-half-even.
    port, _ = libs
    W, H = shape
    for scale0, nlevels in [(1.05, 64), (1.02, 128), (1.05, 15), (1.0, 64), (1.3, 3)]:
        sc, lw, lh = port.levels(W, H, (64, 128), scale0, nlevels)
        exp = []
        s = 1.0
        for _ in range(nlevels):
            exp.append(s)
            if round(W / s) < 64 or round(H / s) < 128 or scale0 <= 1:
                break
            s *= scale0
        else:
            pass
        n = max(len(exp) - 1 if (round(W / exp[-1]) < 64 or round(H / exp[-1]) < 128 or scale0 <= 1) else len(exp), 1)
        n = min(n, nlevels)
        assert len(sc) == n
        assert np.array_equal(sc, np.array(exp[:n]))
        assert list(lw) == [round(W / x) for x in exp[:n]] and list(lh) == [round(H / x) for x in exp[:n]]


@pytest.mark.parametrize("cn", [1, 3, 4])
@pytest.mark.parametrize("shape", [(640, 480), (321, 243), (64, 128), (200, 150)])
def test_resize_linear_bit_exact(libs, cn, shape):
    port, ref = libs
    W, H = shape
    img = ref.rng_fill(1000 + cn, W, H, cn)
    sc, lw, lh = port.levels(W, H, (16, 16), 1.05, 64)
    sizes = set(zip(lw.tolist(), lh.tolist())) | {(W // 2, H // 2), (W // 3, H // 3), (W - 1, H - 1), (W, H),
                                                  (W // 4, H // 4), (17, 9), (W * 2, H * 2), (W + 3, H + 5)}
    for dw, dh in sorted(sizes):
        if dw < 1 or dh < 1:
            continue
        a = port.resize(img, dw, dh)
        b = ref.resize(img, dw, dh)
        assert np.array_equal(a, b), (dw, dh, int(np.abs(a.astype(int) - b).max()))


def test_resize_exact_half_is_area(libs):
    port, ref = libs
    img = ref.rng_fill(7, 1920 // 4, 1080 // 4, 4)
    a = port.resize(img, 240, 135)
    assert np.array_equal(a, ref.resize(img, 240, 135))
    s = img.astype(np.int32)
    exp = (s[0::2, 0::2] + s[0::2, 1::2] + s[1::2, 0::2] + s[1::2, 1::2] + 2) >> 2
    assert np.array_equal(a, exp.astype(np.uint8))


@pytest.mark.parametrize("cn", [1, 3])
@pytest.mark.parametrize("k", [1, 2, 3])
def test_gradient_equals_reference_strict_inputs(libs, cn, k):
    """Inputs as the reference's Strict test builds them (test_cascadeandhog.cpp:1325-1335)."""
    port, ref = libs
    img = ref.rng_fill(0x1234 + k * 7 + cn, 64 * k, 128 * k, cn)
    g0, q0 = port.gradient(img)
    g1, q1 = ref.gradient(img)
    assert np.array_equal(q0, q1)
    assert np.array_equal(g0, g1)


@pytest.mark.parametrize("shape", [(67, 129), (130, 141), (201, 135)])
def test_gradient_tail_columns_and_bgra(libs, shape):
    """width % 4 != 0 exercises the scalar-tail tie-break (hog.cpp:449-479); BGRA == BGR (alpha ignored)."""
    port, ref = libs
    W, H = shape
    bgr = synthetic_frame(5, W, H, 3, blur=False)
    bgr[:, :, 1] = bgr[:, :, 0]          # force channel ties
    g0, q0 = port.gradient(bgr)
    g1, q1 = ref.gradient(bgr)
    assert np.array_equal(q0, q1) and np.array_equal(g0, g1)
    bgra = np.concatenate([bgr, np.full((H, W, 1), 77, np.uint8)], axis=2)
    g2, q2 = port.gradient(bgra)
    assert np.array_equal(q0, q2) and np.array_equal(g0, g2)


@pytest.mark.parametrize("cn", [1, 3])
def test_descriptors_equal_reference(libs, cn):
    port, ref = libs
    img = ref.rng_fill(99 + cn, 64 * 3, 128 * 2, cn)
    for ws in [(8, 8), (64, 128), (16, 8)]:
        a = port.compute(img, ws)
        b = ref.compute(img, ws)
        assert a.shape == b.shape and a.shape[0] == parts(192, 64, ws[0]) * parts(256, 128, ws[1])
        assert np.array_equal(a, b), float(np.abs(a - b).max())


def test_descriptors_other_geometries(libs):
    port, ref = libs
    img = synthetic_frame(3, 160, 200, 1)
    for g in [dict(DEFAULT_GEOM, win=(48, 96)),
              dict(DEFAULT_GEOM, win=(64, 128), gamma=False),
              # NB l2hys is NOT varied here: the reference's SSE2 normaliser keeps the threshold of the first
              # call of the process in a function-local static (objdetect/src/hog.cpp:1168), a known quirk
              dict(DEFAULT_GEOM, win=(64, 128), win_sigma=3.0)]:
        a = port.compute(img, (8, 8), g)
        b = ref.compute(img, (8, 8), g)
        assert np.array_equal(a, b)


@pytest.mark.parametrize("cn", [1, 4])
def test_scores_equal_reference_detect(libs, cn):
    port, ref = libs
    det = ref.people_detector(0)
    img = synthetic_frame(11, 320, 272, cn, silhouettes=3)
    sc = port.scores(img, det)
    xy, wt = ref.detect(img, det, hit_threshold=-1e30)
    assert len(xy) == sc.size
    assert np.array_equal(xy[:, 0].reshape(sc.shape) // 8, np.tile(np.arange(sc.shape[1]), (sc.shape[0], 1)))
    assert np.array_equal(wt.reshape(sc.shape), sc)
    # thresholded hit list, row-major order
    thr = float(np.median(sc))
    xy2, wt2 = ref.detect(img, det, hit_threshold=thr)
    ys, xs = np.nonzero(sc >= thr)
    assert np.array_equal(xy2, np.stack([xs * 8, ys * 8], 1))


def test_group_rectangles_equal_reference(libs):
    port, ref = libs
    rng = np.random.RandomState(0)
    for n, thr in [(0, 2), (1, 1), (50, 1), (300, 2), (1000, 2), (400, 0), (400, 5)]:
        centers = rng.randint(0, 600, size=(max(n // 8, 1), 2))
        idx = rng.randint(0, len(centers), size=n)
        wh = rng.choice([64, 67, 71, 74, 78, 128], size=n)
        r = np.stack([centers[idx, 0] + rng.randint(-9, 10, n), centers[idx, 1] + rng.randint(-9, 10, n),
                      wh, wh * 2], 1).astype(np.int32).reshape(-1, 4)
        a = port.group_rectangles(r, thr)
        b = ref.group_rectangles(r, thr)
        assert np.array_equal(a, b), (n, thr)


def test_detect_multiscale_vs_reference_pieces(libs):
    """The oracle's detectMultiScale == reference resize + reference detect + reference groupRectangles
    composed with cv::cuda::HOG's host logic (cudaobjdetect/src/hog.cpp:415-488)."""
    port, ref = libs
    det = ref.people_detector(0)
    img = synthetic_frame(21, 400, 300, 4, silhouettes=4)
    hit = -2.0
    sc, lw, lh = port.levels(400, 300, (64, 128), 1.05, 64)
    cand = []
    for s, w, h in zip(sc, lw, lh):
        lvl = img if (w, h) == (400, 300) else ref.resize(img, int(w), int(h))
        xy, _ = ref.detect(lvl, det, hit_threshold=hit)
        for x, y in xy:
            cand.append([round(x * s), round(y * s), round(64 * s), round(128 * s)])
    cand = np.array(cand, np.int32).reshape(-1, 4)
    exp = ref.group_rectangles(cand, 2)
    got, _ = port.detect_multiscale(img, det, hit_threshold=hit, group_threshold=2)
    assert len(cand) > 30
    assert np.array_equal(got, exp)
    raw, conf = port.detect_multiscale(img, det, hit_threshold=hit, group_threshold=0)
    assert np.array_equal(raw, cand) and np.all(conf >= hit)


def test_random_shapes_fuzz():
    """Random sizes (odd widths, scalar-tail columns, exact halves, up- and down-scaling), blurred or raw noise, both
    channel counts: pyramid levels and window descriptors of the restatement are BIT-equal to the unmodified reference."""
    ref, port = Ref(), Port()
    for seed in range(24):
        rng = np.random.RandomState(seed)
        W = int(rng.randint(64, 420)); H = int(rng.randint(128, 360)); cn = int(rng.choice([1, 4]))
        img = synthetic_frame(seed, W, H, cn, blur=bool(rng.rand() < 0.7), silhouettes=int(rng.randint(0, 3)))
        half = rng.rand() < 0.2
        dw = W // 2 if half else int(rng.randint(64, W + 40)); dh = H // 2 if half else int(rng.randint(64, H + 40))
        assert np.array_equal(port.resize(img, dw, dh), ref.resize(img, dw, dh)), (seed, W, H, cn, dw, dh)
        bgr = np.ascontiguousarray(img[:, :, :3]) if cn == 4 else img
        assert np.array_equal(port.compute(img, win_stride=(8, 8)), ref.compute(bgr, win_stride=(8, 8))), (seed, W, H, cn)
