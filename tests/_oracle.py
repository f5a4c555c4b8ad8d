"""ctypes wrappers over the CHECKERS (test infrastructure, never the product path):

* ``Port``  -- oracle/libhogoracle.so, the plain-C restatement (oracle/hog_oracle.c)
* ``Ref``   -- oracle/_ref/libcvref.so, the unmodified reference CPU implementation built by
               oracle/Makefile from /root/reference (present when it was built in the container;
               the built .so travels to the GPU box)

Both expose the same numpy-level API so tests can run one against the other.
"""
import ctypes as C
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PORT_SO = os.path.join(ROOT, "oracle", "libhogoracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libcvref.so")

DEFAULT_GEOM = dict(win=(64, 128), block=(16, 16), stride=(8, 8), cell=(8, 8), nbins=9,
                    win_sigma=-1.0, l2hys=0.2, gamma=True)


def _vp(a):
    return a.ctypes.data_as(C.c_void_p)


def _img(img):
    img = np.ascontiguousarray(img, dtype=np.uint8)
    if img.ndim == 2:
        img = img[:, :, None]
    h, w, cn = img.shape
    return img, w, h, cn, C.c_long(w * cn)


class _Geom(C.Structure):
    _fields_ = [("win_w", C.c_int), ("win_h", C.c_int), ("block_w", C.c_int), ("block_h", C.c_int),
                ("stride_w", C.c_int), ("stride_h", C.c_int), ("cell_w", C.c_int), ("cell_h", C.c_int),
                ("nbins", C.c_int), ("win_sigma", C.c_double), ("l2hys", C.c_double), ("gamma", C.c_int)]


def _geom_struct(g):
    return _Geom(g["win"][0], g["win"][1], g["block"][0], g["block"][1], g["stride"][0], g["stride"][1],
                 g["cell"][0], g["cell"][1], g["nbins"], g["win_sigma"], g["l2hys"], int(g["gamma"]))


def desc_size(g):
    nbw = (g["win"][0] - g["block"][0]) // g["stride"][0] + 1
    nbh = (g["win"][1] - g["block"][1]) // g["stride"][1] + 1
    return nbw * nbh * hist_size(g)


def hist_size(g):
    return (g["block"][0] // g["cell"][0]) * (g["block"][1] // g["cell"][1]) * g["nbins"]


def parts(size, part, stride):
    return (size - part) // stride + 1 if size >= part else 0


class Port:
    """oracle/libhogoracle.so"""
    kind = "port"

    def __init__(self):
        if not os.path.exists(PORT_SO) or os.path.getmtime(PORT_SO) < os.path.getmtime(
                os.path.join(ROOT, "oracle", "hog_oracle.c")):
            subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "port"])
        self.L = L = C.CDLL(PORT_SO)
        for f in ("hogo_block_hists", "hogo_compute", "hogo_scores", "hogo_group_rectangles",
                  "hogo_detect_multiscale"):
            getattr(L, f).restype = C.c_long

    def levels(self, W, H, win=(64, 128), scale0=1.05, nlevels=64):
        cap = 1024
        sc = np.zeros(cap, np.float64); lw = np.zeros(cap, np.int32); lh = np.zeros(cap, np.int32)
        n = self.L.hogo_levels(W, H, win[0], win[1], C.c_double(scale0), nlevels, _vp(sc), _vp(lw), _vp(lh), cap)
        return sc[:n].copy(), lw[:n].copy(), lh[:n].copy()

    def resize(self, img, dw, dh):
        img, w, h, cn, pitch = _img(img)
        out = np.zeros((dh, dw, cn), np.uint8)
        self.L.hogo_resize_linear(_vp(img), w, h, cn, pitch, _vp(out), dw, dh, C.c_long(dw * cn))
        return out

    def gradient(self, img, g=DEFAULT_GEOM):
        img, w, h, cn, pitch = _img(img)
        grad = np.zeros((h, w, 2), np.float32); qa = np.zeros((h, w, 2), np.uint8)
        gs = _geom_struct(g)
        self.L.hogo_compute_gradient(C.byref(gs), _vp(img), w, h, cn, pitch, _vp(grad), _vp(qa))
        return grad, qa

    def block_hists(self, img, g=DEFAULT_GEOM):
        img, w, h, cn, pitch = _img(img)
        nbx = parts(w, g["block"][0], g["stride"][0]); nby = parts(h, g["block"][1], g["stride"][1])
        out = np.zeros((max(nby, 0), max(nbx, 0), hist_size(g)), np.float32)
        gs = _geom_struct(g); bx = C.c_int(); by = C.c_int()
        self.L.hogo_block_hists(C.byref(gs), _vp(img), w, h, cn, pitch, _vp(out), C.byref(bx), C.byref(by))
        return out

    def compute(self, img, win_stride=(8, 8), g=DEFAULT_GEOM):
        img, w, h, cn, pitch = _img(img)
        wx = parts(w, g["win"][0], win_stride[0]); wy = parts(h, g["win"][1], win_stride[1])
        out = np.zeros((max(wx * wy, 0), desc_size(g)), np.float32)
        gs = _geom_struct(g)
        self.L.hogo_compute(C.byref(gs), _vp(img), w, h, cn, pitch, win_stride[0], win_stride[1], _vp(out))
        return out

    def scores(self, img, svm, win_stride=(8, 8), g=DEFAULT_GEOM):
        img, w, h, cn, pitch = _img(img)
        svm = np.ascontiguousarray(svm, np.float32)
        wx = parts(w, g["win"][0], win_stride[0]); wy = parts(h, g["win"][1], win_stride[1])
        out = np.zeros((max(wy, 0), max(wx, 0)), np.float64)
        gs = _geom_struct(g); a = C.c_int(); b = C.c_int()
        self.L.hogo_scores(C.byref(gs), _vp(img), w, h, cn, pitch, _vp(svm), svm.size, win_stride[0],
                           win_stride[1], _vp(out), C.byref(a), C.byref(b))
        return out

    def group_rectangles(self, rects, thr, eps=0.2):
        r = np.ascontiguousarray(rects, np.int32).reshape(-1, 4).copy()
        n = self.L.hogo_group_rectangles(_vp(r), C.c_long(len(r)), thr, C.c_double(eps))
        return r[:n].copy()

    def detect_multiscale(self, img, svm, hit_threshold=0.0, win_stride=(8, 8), scale0=1.05, nlevels=64,
                          group_threshold=2, g=DEFAULT_GEOM, capacity=2_000_000):
        """cv::cuda::HOG::detectMultiScale semantics on the INTER_LINEAR pyramid."""
        img, w, h, cn, pitch = _img(img)
        svm = np.ascontiguousarray(svm, np.float32)
        rects = np.zeros((capacity, 4), np.int32); conf = np.zeros(capacity, np.float64)
        gs = _geom_struct(g)
        n = self.L.hogo_detect_multiscale(C.byref(gs), _vp(img), w, h, cn, pitch, _vp(svm), svm.size,
                                          C.c_double(hit_threshold), win_stride[0], win_stride[1],
                                          C.c_double(scale0), nlevels, group_threshold, _vp(rects), _vp(conf),
                                          C.c_long(capacity))
        if n < 0:
            raise RuntimeError("oracle capacity")
        return rects[:n].copy(), conf[:n].copy()


class Ref:
    """oracle/_ref/libcvref.so -- the unmodified reference."""
    kind = "reference"

    @staticmethod
    def available():
        return os.path.exists(REF_SO)

    def __init__(self):
        self.L = L = C.CDLL(REF_SO)
        L.cvref_last_error.restype = C.c_char_p
        L.cvref_version.restype = C.c_char_p
        for f in ("cvref_compute", "cvref_detect", "cvref_detect_multiscale", "cvref_group_rectangles",
                  "cvref_people_detector"):
            getattr(L, f).restype = C.c_long

    def _chk(self, r):
        if r < 0:
            raise RuntimeError(self.L.cvref_last_error().decode())
        return r

    @staticmethod
    def _g(g):
        return ((C.c_int * 9)(g["win"][0], g["win"][1], g["block"][0], g["block"][1], g["stride"][0],
                              g["stride"][1], g["cell"][0], g["cell"][1], g["nbins"]),
                C.c_double(g["win_sigma"]), C.c_double(g["l2hys"]), int(g["gamma"]))

    def version(self):
        return self.L.cvref_version().decode()

    def set_num_threads(self, n):
        self.L.cvref_set_num_threads(n)

    def get_num_threads(self):
        return self.L.cvref_get_num_threads()

    def people_detector(self, which=0):
        out = np.zeros(4096, np.float32)
        n = self._chk(self.L.cvref_people_detector(which, _vp(out), C.c_long(out.size)))
        return out[:n].copy()

    def hog_save(self, path, detector, objname=None, g=DEFAULT_GEOM, nlevels=64):
        """cv::HOGDescriptor::save of the unmodified reference (objdetect/src/hog.cpp:205-209)."""
        geom, ws, l2, gamma = self._g(g)
        d = None if detector is None else np.ascontiguousarray(detector, np.float32)
        self.L.cvref_hog_save.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(C.c_int), C.c_double, C.c_double, C.c_int,
                                          C.c_int, C.c_void_p, C.c_long]
        self._chk(self.L.cvref_hog_save(os.fsencode(path), objname.encode() if objname else None, geom, ws, l2, gamma,
                                        nlevels, None if d is None else d.ctypes.data, 0 if d is None else d.size))

    def hog_load(self, path, objname=None):
        """cv::HOGDescriptor::load of the unmodified reference (hog.cpp:198-203)."""
        geom = (C.c_int * 9)(); ws = C.c_double(); l2 = C.c_double(); gamma = C.c_int(); nl = C.c_int()
        det = np.zeros(1 << 16, np.float32)
        self.L.cvref_hog_load.restype = C.c_long
        self.L.cvref_hog_load.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_double),
                                          C.POINTER(C.c_double), C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_void_p, C.c_long]
        n = self._chk(self.L.cvref_hog_load(os.fsencode(path), objname.encode() if objname else None, geom, C.byref(ws),
                                            C.byref(l2), C.byref(gamma), C.byref(nl), det.ctypes.data, det.size))
        return {"geom": list(geom), "win_sigma": ws.value, "l2hys_threshold": l2.value,
                "gamma_correction": bool(gamma.value), "nlevels": nl.value, "detector": det[:n].copy()}

    def rng_fill(self, seed, w, h, cn):
        img = np.zeros((h, w, cn), np.uint8)
        self._chk(self.L.cvref_rng_fill(C.c_ulonglong(seed), _vp(img), w, h, cn, C.c_long(w * cn)))
        return img

    def resize(self, img, dw, dh, exact=False):
        img, w, h, cn, pitch = _img(img)
        out = np.zeros((dh, dw, cn), np.uint8)
        self._chk(self.L.cvref_resize(_vp(img), w, h, cn, pitch, _vp(out), dw, dh, C.c_long(dw * cn), int(exact)))
        return out

    def gradient(self, img, g=DEFAULT_GEOM):
        img, w, h, cn, pitch = _img(img)
        grad = np.zeros((h, w, 2), np.float32); qa = np.zeros((h, w, 2), np.uint8)
        self._chk(self.L.cvref_compute_gradient(*self._g(g), _vp(img), w, h, cn, pitch, _vp(grad), _vp(qa)))
        return grad, qa

    def compute(self, img, win_stride=(8, 8), g=DEFAULT_GEOM):
        img, w, h, cn, pitch = _img(img)
        wx = parts(w, g["win"][0], win_stride[0]); wy = parts(h, g["win"][1], win_stride[1])
        out = np.zeros((max(wx * wy, 0), desc_size(g)), np.float32)
        n = self._chk(self.L.cvref_compute(*self._g(g), _vp(img), w, h, cn, pitch, win_stride[0], win_stride[1],
                                           _vp(out), C.c_long(out.size)))
        assert n == out.size, (n, out.size)
        return out

    def detect(self, img, svm, hit_threshold=0.0, win_stride=(8, 8), g=DEFAULT_GEOM):
        img, w, h, cn, pitch = _img(img)
        svm = np.ascontiguousarray(svm, np.float32)
        cap = max(parts(w, g["win"][0], win_stride[0]) * parts(h, g["win"][1], win_stride[1]), 1)
        xy = np.zeros((cap, 2), np.int32); wt = np.zeros(cap, np.float64)
        n = self._chk(self.L.cvref_detect(*self._g(g), _vp(svm), svm.size, _vp(img), w, h, cn, pitch,
                                          C.c_double(hit_threshold), win_stride[0], win_stride[1],
                                          _vp(xy), _vp(wt), C.c_long(cap)))
        return xy[:n].copy(), wt[:n].copy()

    def detect_multiscale_stock(self, img, svm, hit_threshold=0.0, win_stride=(8, 8), scale0=1.05, nlevels=64,
                                final_threshold=2.0, g=DEFAULT_GEOM, capacity=2_000_000):
        """Stock cv::HOGDescriptor::detectMultiScale (INTER_LINEAR_EXACT pyramid) -- the timed CPU baseline."""
        img, w, h, cn, pitch = _img(img)
        svm = np.ascontiguousarray(svm, np.float32)
        rects = np.zeros((capacity, 4), np.int32); wt = np.zeros(capacity, np.float64)
        n = self._chk(self.L.cvref_detect_multiscale(*self._g(g), nlevels, _vp(svm), svm.size, _vp(img), w, h, cn,
                                                     pitch, C.c_double(hit_threshold), win_stride[0], win_stride[1],
                                                     C.c_double(scale0), C.c_double(final_threshold),
                                                     _vp(rects), _vp(wt), C.c_long(capacity)))
        return rects[:n].copy(), wt[:n].copy()

    def group_rectangles(self, rects, thr, eps=0.2):
        r = np.ascontiguousarray(rects, np.int32).reshape(-1, 4).copy()
        n = self._chk(self.L.cvref_group_rectangles(_vp(r), C.c_long(len(r)), thr, C.c_double(eps)))
        return r[:n].copy()

    def group_rectangles_meanshift(self, rects, weights, scales, detect_threshold=0.0, win=(64, 128)):
        """cv::groupRectangles_meanshift of the unmodified reference (cascadedetect.cpp:387-394)."""
        r = np.ascontiguousarray(rects, np.int32).reshape(-1, 4).copy()
        w = np.ascontiguousarray(weights, np.float64).copy(); sc = np.ascontiguousarray(scales, np.float64)
        self.L.cvref_group_rectangles_meanshift.restype = C.c_long
        n = self._chk(self.L.cvref_group_rectangles_meanshift(_vp(r), _vp(w), _vp(sc), C.c_long(len(r)),
                                                              C.c_double(detect_threshold), win[0], win[1]))
        return r[:n].copy(), w[:n].copy()

    def detect_roi(self, img, svm, locations, hit_threshold=0.0, padding=(8, 8), g=DEFAULT_GEOM):
        """HOGDescriptor::detectROI of the unmodified reference (objdetect/src/hog.cpp:3549-3660)
        -> (found (M,2) int32, confidences (len(locations),) float64)."""
        img, w, h, cn, pitch = _img(img)
        svm = np.ascontiguousarray(svm, np.float32)
        loc = np.ascontiguousarray(locations, np.int32).reshape(-1, 2)
        found = np.zeros((max(len(loc), 1), 2), np.int32); conf = np.zeros(max(len(loc), 1), np.float64)
        self.L.cvref_detect_roi.restype = C.c_long
        n = self._chk(self.L.cvref_detect_roi(*self._g(g), _vp(svm), svm.size, _vp(img), w, h, cn, pitch, _vp(loc),
                                              C.c_long(len(loc)), C.c_double(hit_threshold), padding[0], padding[1],
                                              _vp(found), _vp(conf)))
        return found[:n].copy(), conf[:len(loc)].copy()

    def detect_multiscale_roi(self, img, svm, rois, hit_threshold=0.0, group_threshold=0, g=DEFAULT_GEOM, capacity=1 << 20):
        """HOGDescriptor::detectMultiScaleROI, unmodified (hog.cpp:3662-3677).  rois: list of (scale, (K,2) locations)
        -> (rects (M,4), [confidences per roi])."""
        img, w, h, cn, pitch = _img(img)
        svm = np.ascontiguousarray(svm, np.float32)
        scales = np.array([r[0] for r in rois], np.float64)
        locs = [np.asarray(r[1], np.int32).reshape(-1, 2) for r in rois]
        off = np.zeros(len(rois) + 1, np.int64); off[1:] = np.cumsum([len(l) for l in locs])
        allloc = np.ascontiguousarray(np.concatenate(locs) if locs else np.zeros((0, 2), np.int32))
        rects = np.zeros((capacity, 4), np.int32); conf = np.zeros(max(int(off[-1]), 1), np.float64)
        self.L.cvref_detect_multiscale_roi.restype = C.c_long
        n = self._chk(self.L.cvref_detect_multiscale_roi(*self._g(g), _vp(svm), svm.size, _vp(img), w, h, cn, pitch, _vp(scales),
                                                         _vp(off), C.c_long(len(rois)), _vp(allloc), C.c_double(hit_threshold),
                                                         group_threshold, _vp(rects), C.c_long(capacity), _vp(conf)))
        return rects[:n].copy(), [conf[off[i]:off[i + 1]].copy() for i in range(len(rois))]


def synthetic_frame(seed, w, h, cn, blur=True, silhouettes=0):
    """Deterministic synthetic frame, numpy only (identical bytes on every machine).

    Uniform noise smoothed by a separable 5-tap binomial filter (natural-ish gradients), plus optional
    bright vertical ellipses so that some windows score high.  Channel 3 of a 4-channel frame is 255."""
    rng = np.random.RandomState(seed)
    c = 3 if cn == 4 else cn
    img = rng.randint(0, 256, size=(h, w, c)).astype(np.float32)
    if blur:
        k = np.array([1, 4, 6, 4, 1], np.float32) / 16
        for _ in range(2):
            p = np.pad(img, ((2, 2), (0, 0), (0, 0)), mode="reflect")
            img = sum(k[i] * p[i:i + h] for i in range(5))
            p = np.pad(img, ((0, 0), (2, 2), (0, 0)), mode="reflect")
            img = sum(k[i] * p[:, i:i + w] for i in range(5))
        img = (img - img.mean()) * 3 + 128
    for s in range(silhouettes):
        hh = rng.randint(48, max(49, min(200, h // 2)))
        ww = max(hh // 3, 8)
        cy = rng.randint(hh // 2, max(hh // 2 + 1, h - hh // 2)); cx = rng.randint(ww // 2, max(ww // 2 + 1, w - ww // 2))
        yy, xx = np.ogrid[:h, :w]
        m = ((yy - cy) / (hh / 2.0)) ** 2 + ((xx - cx) / (ww / 2.0)) ** 2 <= 1
        img[m] = img[m] * 0.2 + 200
    img = np.clip(np.rint(img), 0, 255).astype(np.uint8)
    if cn == 4:
        img = np.concatenate([img, np.full((h, w, 1), 255, np.uint8)], axis=2)
    return np.ascontiguousarray(img)


SCORE_TOL = 1e-3


def detections_match(hog, port, img, det, hit, group, scale0=1.05, nlevels=64, slot=0, g=DEFAULT_GEOM, win_stride=(8, 8)):
    """The north-star bar for detectMultiScale: the GPU's grouped set equals the oracle's, except for windows whose
    score lies within SCORE_TOL of the hit threshold.  When the two differ, (1) the GPU's UNGROUPED candidate set must
    lie between the oracle's sets at hit +/- SCORE_TOL (only near-threshold windows may differ), and (2) the GPU's
    grouped output must be exactly what the oracle's groupRectangles makes of the GPU's own candidate list -- so the
    grouped output is verified in both cases.  -> (ok, got, exp); leaves hit/group thresholds set on `hog`."""
    hog.setHitThreshold(hit); hog.setGroupThreshold(group); hog.setScaleFactor(scale0); hog.setNumLevels(nlevels)
    got = hog.detectMultiScale(img, slot=slot)
    kw = dict(win_stride=win_stride, scale0=scale0, nlevels=nlevels, g=g)
    exp, _ = port.detect_multiscale(img, det, hit_threshold=hit, group_threshold=group, **kw)
    if np.array_equal(got, exp):
        return True, got, exp
    lo, _ = port.detect_multiscale(img, det, hit_threshold=hit - SCORE_TOL, group_threshold=0, **kw)
    hi, _ = port.detect_multiscale(img, det, hit_threshold=hit + SCORE_TOL, group_threshold=0, **kw)
    hog.setGroupThreshold(0)
    raw = hog.detectMultiScale(img, slot=slot)
    hog.setGroupThreshold(group)
    s_lo = set(map(tuple, lo)); s_hi = set(map(tuple, hi)); s_raw = set(map(tuple, raw))
    ok = s_hi <= s_raw <= s_lo and len(s_lo - s_hi) > 0
    regrouped = port.group_rectangles(raw, group) if group > 0 else raw
    ok = ok and np.array_equal(regrouped, got)
    return ok, got, exp
