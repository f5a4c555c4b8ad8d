"""CPU: the C oracle (oracle/hog_oracle.c) against the committed golden vectors, which are outputs of the
UNMODIFIED reference (tests/golden/make_golden.py).  Runs anywhere -- /root/reference is not needed."""
import os

import numpy as np
import pytest

from _oracle import Port

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "hog_golden.npz"))
RESIZE4 = [(91, 57), (80, 50), (48, 30), (33, 21), (17, 9)]
RESIZE1 = [(92, 58), (77, 48), (41, 26)]


@pytest.fixture(scope="module")
def port():
    return Port()


def bgr(img):
    return np.ascontiguousarray(img[:, :, :3])


def test_resize_golden(port):
    for i, (dw, dh) in enumerate(RESIZE4):
        assert np.array_equal(port.resize(G["resize_src4"], dw, dh), G["resize_dst4_%d" % i]), (dw, dh)
    for i, (dw, dh) in enumerate(RESIZE1):
        assert np.array_equal(port.resize(G["resize_src1"], dw, dh), G["resize_dst1_%d" % i]), (dw, dh)


def test_descriptor_golden(port):
    assert np.array_equal(port.compute(G["desc_img1"]), G["desc_1"])
    assert np.array_equal(port.compute(G["desc_img4"]), G["desc_4"])          # the port ignores channel 3
    assert np.array_equal(port.compute(bgr(G["desc_img4"])), G["desc_4"])
    assert np.array_equal(port.compute(G["desc_img4_tail"]), G["desc_4_tail"])


def test_scores_golden(port):
    sc = port.scores(G["score_img4"], G["people_detector_64x128"])
    xy = G["score_xy"]
    assert sc.size == len(xy)
    got = sc[xy[:, 1] // 8, xy[:, 0] // 8]
    assert np.array_equal(got, G["score_w"])


def test_group_rectangles_golden(port):
    assert np.array_equal(port.group_rectangles(G["group_in"], 2), G["group_out_thr2"])
    assert np.array_equal(port.group_rectangles(G["group_in"], 1), G["group_out_thr1"])


def test_detect_multiscale_golden(port):
    hit = float(G["ms_hit_threshold"][0])
    sc, lw, lh = port.levels(320, 272)
    assert np.array_equal(sc, G["ms_scales"])
    cand, _ = port.detect_multiscale(G["ms_img4"], G["people_detector_64x128"], hit_threshold=hit, group_threshold=0)
    assert np.array_equal(cand, G["ms_candidates"])
    grp, _ = port.detect_multiscale(G["ms_img4"], G["people_detector_64x128"], hit_threshold=hit, group_threshold=2)
    assert sorted(map(tuple, grp)) == sorted(map(tuple, G["ms_grouped_thr2"]))
