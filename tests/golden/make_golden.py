"""Generates tests/golden/hog_golden.npz from the UNMODIFIED reference (oracle/_ref/libcvref.so, built from
/root/reference by oracle/Makefile).  Run in the build container:

    python tests/golden/make_golden.py

The reference ships no usable golden files for this path (its GPU golden test is disabled and its data lives in
opencv_extra, which is absent -- SURVEY.md section 4), so these vectors ARE outputs of the reference itself on
small seeded inputs.  They travel to the GPU box, where /root/reference does not exist, and pin both the C
oracle (CPU tests) and the CUDA path (GPU tests).  Inputs are stored too, so nothing depends on an RNG.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from _oracle import Ref, synthetic_frame  # noqa: E402


def bgr(img):
    return np.ascontiguousarray(img[:, :, :3]) if img.ndim == 3 and img.shape[2] == 4 else img


def main():
    ref = Ref()
    assert ref.version() == "3.4.7"
    out = {}
    det = ref.people_detector(0)
    out["people_detector_64x128"] = det
    out["people_detector_48x96"] = ref.people_detector(1)

    # ---- cv::resize INTER_LINEAR (bit-exact target), 8UC4 and 8UC1, incl. the exact-half INTER_AREA switch
    src4 = synthetic_frame(11, 96, 60, 4, blur=False)
    src1 = synthetic_frame(12, 97, 61, 1, blur=False)
    out["resize_src4"] = src4
    out["resize_src1"] = src1
    for i, (dw, dh) in enumerate([(91, 57), (80, 50), (48, 30), (33, 21), (17, 9)]):
        out["resize_dst4_%d" % i] = ref.resize(src4, dw, dh)
    for i, (dw, dh) in enumerate([(92, 58), (77, 48), (41, 26)]):
        out["resize_dst1_%d" % i] = ref.resize(src1, dw, dh)

    # ---- cv::HOGDescriptor::compute: one window, 8UC1 and colour (BGR view of the BGRA bytes)
    win1 = synthetic_frame(21, 64, 128, 1)
    win4 = synthetic_frame(22, 64, 128, 4)
    out["desc_img1"] = win1
    out["desc_img4"] = win4
    out["desc_1"] = ref.compute(win1, (8, 8))
    out["desc_4"] = ref.compute(bgr(win4), (8, 8))
    # a width with a scalar tail (w % 4 != 0) and several windows
    tail4 = synthetic_frame(23, 75, 136, 4, blur=False)
    out["desc_img4_tail"] = tail4
    out["desc_4_tail"] = ref.compute(bgr(tail4), (8, 8))

    # ---- cv::HOGDescriptor::detect with hitThreshold = -inf: score of every window
    sc_img = synthetic_frame(31, 160, 192, 4, silhouettes=2)
    out["score_img4"] = sc_img
    xy, wt = ref.detect(bgr(sc_img), det, hit_threshold=-1e30, win_stride=(8, 8))
    out["score_xy"] = xy
    out["score_w"] = wt

    # ---- cv::groupRectangles(rects, 2, 0.2)
    rng = np.random.RandomState(5)
    centers = rng.randint(0, 300, size=(12, 2))
    rects = []
    for cx, cy in centers:
        for _ in range(rng.randint(1, 7)):
            s = rng.choice([64, 67, 71, 74])
            rects.append([cx + rng.randint(-4, 5), cy + rng.randint(-4, 5), s, 2 * s])
    rects = np.array(rects, np.int32)
    out["group_in"] = rects
    out["group_out_thr2"] = ref.group_rectangles(rects, 2, 0.2)
    out["group_out_thr1"] = ref.group_rectangles(rects, 1, 0.2)

    # ---- the whole path on the mandated pyramid: cv::resize(INTER_LINEAR) levels -> HOGDescriptor::detect per
    #      level -> rect scaling (hog.cpp:473-481) -> free cv::groupRectangles (hog.cpp:486)
    ms = synthetic_frame(41, 320, 272, 4, silhouettes=3)
    out["ms_img4"] = ms
    hit = -1.5
    W, H = 320, 272
    scales = []
    s = 1.0
    for _ in range(64):
        scales.append(s)
        if round(W / s) < 64 or round(H / s) < 128:
            break
        s *= 1.05
    if round(W / scales[-1]) < 64 or round(H / scales[-1]) < 128:
        scales = scales[:-1]
    cand = []
    for s in scales:
        lw, lh = int(round(W / s)), int(round(H / s))          # Python round == cvRound (half to even)
        lvl = ms if (lw, lh) == (W, H) else ref.resize(ms, lw, lh)
        xy, wt = ref.detect(bgr(lvl), det, hit_threshold=hit, win_stride=(8, 8))
        for (x, y) in xy:
            cand.append([int(round(x * s)), int(round(y * s)), int(round(64 * s)), int(round(128 * s))])
    cand = np.array(cand, np.int32).reshape(-1, 4)
    out["ms_hit_threshold"] = np.array([hit])
    out["ms_scales"] = np.array(scales)
    out["ms_candidates"] = cand
    out["ms_grouped_thr2"] = ref.group_rectangles(cand, 2, 0.2)

    path = os.path.join(HERE, "hog_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", len(cand), "candidates,", len(out["ms_grouped_thr2"]), "groups")


if __name__ == "__main__":
    main()
