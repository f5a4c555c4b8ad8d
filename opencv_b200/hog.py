"""ctypes mirror of cv::cuda::HOG (modules/cudaobjdetect/include/opencv2/cudaobjdetect.hpp:78-212) over the
C ABI of include/hogb200.h.  Same method names, argument meaning and error behaviour (errors raise HogError where
the reference throws cv::Exception).  Images are numpy uint8 arrays (host) or torch CUDA uint8 tensors (device),
HxW (8UC1) or HxWx4 (8UC4).

There is no fallback: if libhogb200.so is missing or no CUDA device exists, construction raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def lib_path():
    # HOGB200_LIB: tuning experiments load an alternative build of the same library (tools/build_variant.sh)
    return os.environ.get("HOGB200_LIB") or os.path.join(_HERE, "libhogb200.so")


class HogError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__("hogb200 status %d: %s" % (status, msg))
        self.status = status


class _Geom(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("win_w", "win_h", "block_w", "block_h", "block_stride_w", "block_stride_h",
                                       "cell_w", "cell_h", "nbins")]


class _Image(C.Structure):
    _fields_ = [("data", C.c_void_p), ("pitch", C.c_size_t), ("width", C.c_int), ("height", C.c_int),
                ("channels", C.c_int), ("on_device", C.c_int)]


class _Model(C.Structure):
    _fields_ = [("geom", _Geom), ("win_sigma", C.c_double), ("l2hys_threshold", C.c_double),
                ("gamma_correction", C.c_int), ("nlevels", C.c_int)]


class _Roi(C.Structure):
    _fields_ = [("scale", C.c_double), ("locations", C.c_void_p), ("nlocations", C.c_size_t), ("confidences", C.c_void_p)]


class Timings(C.Structure):
    _fields_ = [("upload_ms", C.c_float), ("resize_ms", C.c_float), ("hist_ms", C.c_float), ("svm_ms", C.c_float),
                ("download_ms", C.c_float), ("total_ms", C.c_float), ("kernel_launches", C.c_int),
                ("pyramid_pixels", C.c_longlong), ("blocks", C.c_longlong), ("windows", C.c_longlong),
                ("candidates", C.c_longlong)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


def load_library():
    """Loads libhogb200.so; raises (never falls back) when it is missing."""
    global _LIB
    if _LIB is not None:
        return _LIB
    p = lib_path()
    if not os.path.exists(p):
        raise HogError(2, "libhogb200.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                          "there is no CPU fallback")
    L = C.CDLL(p)
    L.hogb200_last_error.restype = C.c_char_p
    L.hogb200_version.restype = C.c_char_p
    for f in ("win_sigma", "l2hys_threshold", "hit_threshold", "scale_factor"):
        getattr(L, "hogb200_get_" + f).restype = C.c_double
        getattr(L, "hogb200_set_" + f).argtypes = [C.c_void_p, C.c_double]
    for f in ("win_sigma", "l2hys_threshold", "hit_threshold", "scale_factor", "gamma_correction", "num_levels",
              "group_threshold", "descriptor_format", "descriptor_size", "block_histogram_size"):
        getattr(L, "hogb200_get_" + f).argtypes = [C.c_void_p]
    for f in ("gamma_correction", "num_levels", "group_threshold", "descriptor_format"):
        getattr(L, "hogb200_set_" + f).argtypes = [C.c_void_p, C.c_int]
    L.hogb200_get_descriptor_size.restype = C.c_size_t
    L.hogb200_get_block_histogram_size.restype = C.c_size_t
    L.hogb200_create.argtypes = [C.POINTER(_Geom), C.c_int, C.POINTER(C.c_void_p)]
    L.hogb200_destroy.argtypes = [C.c_void_p]
    L.hogb200_destroy.restype = None
    L.hogb200_set_stream.argtypes = [C.c_void_p, C.c_void_p]
    L.hogb200_set_win_stride.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.hogb200_get_win_stride.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.hogb200_get_win_stride.restype = None
    L.hogb200_set_svm_detector.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
    L.hogb200_default_people_detector.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
    L.hogb200_detect.argtypes = [C.c_void_p, C.POINTER(_Image), C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
    L.hogb200_detect_multiscale.argtypes = L.hogb200_detect.argtypes
    L.hogb200_detect_multiscale_slot.argtypes = [C.c_void_p, C.c_int, C.POINTER(_Image), C.c_void_p, C.c_void_p,
                                                 C.c_size_t, C.POINTER(C.c_size_t)]
    L.hogb200_detect_multiscale_batch.argtypes = [C.c_void_p, C.POINTER(_Image), C.c_int, C.c_int, C.c_void_p,
                                                  C.c_size_t, C.c_void_p]
    L.hogb200_submit_batch.argtypes = [C.c_void_p, C.POINTER(_Image), C.c_int, C.c_int, C.POINTER(C.c_int)]
    L.hogb200_collect_batch.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]
    L.hogb200_compute.argtypes = [C.c_void_p, C.POINTER(_Image), C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_size_t)]
    L.hogb200_levels.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                                 C.POINTER(C.c_size_t)]
    L.hogb200_enumerate_levels.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_void_p,
                                           C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
    L.hogb200_resize.argtypes = [C.c_void_p, C.POINTER(_Image), C.c_int, C.c_int, C.c_void_p]
    L.hogb200_block_histograms.argtypes = [C.c_void_p, C.POINTER(_Image), C.c_void_p, C.c_size_t,
                                           C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.hogb200_block_histograms_noalloc.argtypes = [C.c_void_p, C.POINTER(_Image), C.c_void_p, C.c_size_t,
                                                   C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.hogb200_noalloc_scratch_bytes.restype = C.c_size_t
    L.hogb200_window_scores.argtypes = [C.c_void_p, C.c_int, C.POINTER(_Image), C.c_void_p, C.c_size_t,
                                        C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.hogb200_group_rectangles.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_double, C.POINTER(C.c_size_t)]
    L.hogb200_group_rectangles_device.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_double, C.POINTER(C.c_size_t)]
    L.hogb200_group_rectangles_meanshift.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_double, C.c_int, C.c_int,
                                                     C.POINTER(C.c_size_t)]
    L.hogb200_detect_roi.argtypes = [C.c_void_p, C.POINTER(_Image), C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                     C.c_size_t, C.POINTER(C.c_size_t)]
    L.hogb200_detect_multiscale_roi.argtypes = [C.c_void_p, C.POINTER(_Image), C.POINTER(_Roi), C.c_size_t, C.c_int, C.c_void_p,
                                                C.c_size_t, C.POINTER(C.c_size_t)]
    L.hogb200_batch_group.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
    L.hogb200_model_read.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(_Model), C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
    L.hogb200_model_write.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(_Model), C.c_void_p, C.c_size_t]
    L.hogb200_save.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int]
    L.hogb200_create_from_file.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.c_void_p)]
    L.hogb200_enable_timing.argtypes = [C.c_void_p, C.c_int]
    L.hogb200_last_timings.argtypes = [C.c_void_p, C.POINTER(Timings)]
    _LIB = L
    return L


def _as_image(img, keep):
    """numpy array or torch tensor -> _Image.  `keep` collects objects that must outlive the call."""
    if isinstance(img, np.ndarray):
        if img.dtype != np.uint8:
            raise HogError(1, "image must be uint8")
        if img.ndim == 2:
            h, w = img.shape; cn = 1
        elif img.ndim == 3:
            h, w, cn = img.shape
        else:
            raise HogError(1, "image must be HxW or HxWxC")
        if img.strides[-1] != 1 or (img.ndim == 3 and img.strides[1] != cn):
            img = np.ascontiguousarray(img)
        keep.append(img)
        return _Image(img.ctypes.data, img.strides[0], w, h, cn, 0)
    # torch tensor (duck-typed so that torch stays optional)
    if hasattr(img, "data_ptr") and hasattr(img, "is_cuda"):
        import torch
        if img.dtype != torch.uint8:
            raise HogError(1, "image must be uint8")
        if img.dim() == 2:
            h, w = img.shape; cn = 1
        else:
            h, w, cn = img.shape
        if img.stride(-1) != 1 or (img.dim() == 3 and img.stride(1) != cn):
            img = img.contiguous()
        keep.append(img)
        return _Image(img.data_ptr(), img.stride(0), w, h, cn, 1 if img.is_cuda else 0)
    raise HogError(1, "unsupported image object")


def read_model(filename, objname=None):
    """The "opencv-object-detector-hog" node of an XML / YAML file (cv::HOGDescriptor::load, objdetect/src/hog.cpp:139-203)
    -> dict of parameters + `detector` (float32 array, empty if the file has none).  Host only."""
    L = load_library()
    m = _Model()
    det = np.zeros(1 << 16, np.float32)
    n = C.c_size_t(0)
    rc = L.hogb200_model_read(os.fsencode(filename), objname.encode() if objname else None, C.byref(m),
                              det.ctypes.data, det.size, C.byref(n))
    if rc != 0:
        raise HogError(rc, L.hogb200_last_error().decode())
    g = m.geom
    return {"win_size": (g.win_w, g.win_h), "block_size": (g.block_w, g.block_h),
            "block_stride": (g.block_stride_w, g.block_stride_h), "cell_size": (g.cell_w, g.cell_h), "nbins": g.nbins,
            "win_sigma": m.win_sigma, "l2hys_threshold": m.l2hys_threshold, "gamma_correction": bool(m.gamma_correction),
            "nlevels": m.nlevels, "detector": det[:n.value].copy()}


def write_model(filename, detector=None, objname=None, win_size=(64, 128), block_size=(16, 16), block_stride=(8, 8),
                cell_size=(8, 8), nbins=9, win_sigma=-1.0, l2hys_threshold=0.2, gamma_correction=True, nlevels=64):
    """cv::HOGDescriptor::save (objdetect/src/hog.cpp:176-209), SVMDetector included.  Host only."""
    L = load_library()
    m = _Model(_Geom(win_size[0], win_size[1], block_size[0], block_size[1], block_stride[0], block_stride[1],
                     cell_size[0], cell_size[1], nbins), win_sigma, l2hys_threshold, int(bool(gamma_correction)), nlevels)
    d = None if detector is None else np.ascontiguousarray(detector, np.float32).ravel()
    rc = L.hogb200_model_write(os.fsencode(filename), objname.encode() if objname else None, C.byref(m),
                               d.ctypes.data if d is not None else None, 0 if d is None else d.size)
    if rc != 0:
        raise HogError(rc, L.hogb200_last_error().decode())


class HOG:
    """cv::cuda::HOG look-alike.  Create with HOG.create(...)."""
    DESCR_FORMAT_COL_BY_COL = 0
    DESCR_FORMAT_ROW_BY_ROW = 1

    def __init__(self, win_size=(64, 128), block_size=(16, 16), block_stride=(8, 8), cell_size=(8, 8), nbins=9,
                 device=0):
        self._L = load_library()
        self._h = C.c_void_p()
        g = _Geom(win_size[0], win_size[1], block_size[0], block_size[1], block_stride[0], block_stride[1],
                  cell_size[0], cell_size[1], nbins)
        self._win = tuple(win_size)
        self._ck(self._L.hogb200_create(C.byref(g), device, C.byref(self._h)))

    @classmethod
    def create(cls, win_size=(64, 128), block_size=(16, 16), block_stride=(8, 8), cell_size=(8, 8), nbins=9, device=0):
        return cls(win_size, block_size, block_stride, cell_size, nbins, device)

    @classmethod
    def load(cls, filename, objname=None, device=0):
        """cv::cuda::HOG::load (cudaobjdetect/src/hog.cpp:667-672) as a factory; the file's SVMDetector goes to slot 0."""
        self = cls.__new__(cls)
        self._L = load_library()
        self._h = C.c_void_p()
        self._win = tuple(read_model(filename, objname)["win_size"])
        self._ck(self._L.hogb200_create_from_file(os.fsencode(filename), objname.encode() if objname else None, device,
                                                  C.byref(self._h)))
        return self

    def save(self, filename, objname=None, slot=0):
        """cv::cuda::HOG::save (hog.cpp:674-678), with the detector of `slot` (which the fork's writer leaves out)."""
        self._ck(self._L.hogb200_save(self._h, os.fsencode(filename), objname.encode() if objname else None, slot))

    def __del__(self):
        try:
            if getattr(self, "_h", None) and self._h.value:
                self._L.hogb200_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise HogError(rc, self._L.hogb200_last_error().decode())

    # ---- parameters (cudaobjdetect.hpp:101-142)
    def setWinSigma(self, v): self._ck(self._L.hogb200_set_win_sigma(self._h, v))
    def getWinSigma(self): return self._L.hogb200_get_win_sigma(self._h)
    def setL2HysThreshold(self, v): self._ck(self._L.hogb200_set_l2hys_threshold(self._h, v))
    def getL2HysThreshold(self): return self._L.hogb200_get_l2hys_threshold(self._h)
    def setGammaCorrection(self, v): self._ck(self._L.hogb200_set_gamma_correction(self._h, int(bool(v))))
    def getGammaCorrection(self): return bool(self._L.hogb200_get_gamma_correction(self._h))
    def setNumLevels(self, v): self._ck(self._L.hogb200_set_num_levels(self._h, int(v)))
    def getNumLevels(self): return self._L.hogb200_get_num_levels(self._h)
    def setHitThreshold(self, v): self._ck(self._L.hogb200_set_hit_threshold(self._h, v))
    def getHitThreshold(self): return self._L.hogb200_get_hit_threshold(self._h)
    def setWinStride(self, s): self._ck(self._L.hogb200_set_win_stride(self._h, int(s[0]), int(s[1])))

    def getWinStride(self):
        a, b = C.c_int(), C.c_int()
        self._L.hogb200_get_win_stride(self._h, C.byref(a), C.byref(b))
        return (a.value, b.value)

    def setScaleFactor(self, v): self._ck(self._L.hogb200_set_scale_factor(self._h, v))
    def getScaleFactor(self): return self._L.hogb200_get_scale_factor(self._h)
    def setGroupThreshold(self, v): self._ck(self._L.hogb200_set_group_threshold(self._h, int(v)))
    def getGroupThreshold(self): return self._L.hogb200_get_group_threshold(self._h)
    def setDescriptorFormat(self, v): self._ck(self._L.hogb200_set_descriptor_format(self._h, int(v)))
    def getDescriptorFormat(self): return self._L.hogb200_get_descriptor_format(self._h)
    def getDescriptorSize(self): return self._L.hogb200_get_descriptor_size(self._h)
    def getBlockHistogramSize(self): return self._L.hogb200_get_block_histogram_size(self._h)

    def setStream(self, cuda_stream_ptr): self._ck(self._L.hogb200_set_stream(self._h, cuda_stream_ptr))
    def enableTiming(self, on=True): self._ck(self._L.hogb200_enable_timing(self._h, int(on)))

    def lastTimings(self):
        t = Timings()
        self._ck(self._L.hogb200_last_timings(self._h, C.byref(t)))
        return t.as_dict()

    # ---- detectors
    def setSVMDetector(self, detector, slot=0):
        d = np.ascontiguousarray(np.asarray(detector, dtype=np.float32).reshape(-1))
        self._ck(self._L.hogb200_set_svm_detector(self._h, slot, d.ctypes.data, d.size))

    def getDefaultPeopleDetector(self):
        p, n = C.c_void_p(), C.c_size_t()
        self._ck(self._L.hogb200_default_people_detector(self._win[0], self._win[1], C.byref(p), C.byref(n)))
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(n.value,)).copy()

    # ---- work calls
    def detect(self, img, confidences=False):
        """-> hits (N,2) int32 [, confidences (N,) float64]"""
        keep = []; im = _as_image(img, keep)
        cap = 1 << 14
        while True:
            pts = np.zeros((cap, 2), np.int32); conf = np.zeros(cap, np.float64); n = C.c_size_t()
            rc = self._L.hogb200_detect(self._h, C.byref(im), pts.ctypes.data, conf.ctypes.data if confidences else None,
                                        cap, C.byref(n))
            if rc == 3:
                cap = n.value; continue
            self._ck(rc)
            return (pts[:n.value].copy(), conf[:n.value].copy()) if confidences else pts[:n.value].copy()

    def detectMultiScale(self, img, confidences=False, slot=0):
        """-> rects (N,4) int32 (x,y,w,h) [, confidences]"""
        keep = []; im = _as_image(img, keep)
        cap = 1 << 14
        while True:
            r = np.zeros((cap, 4), np.int32); conf = np.zeros(cap, np.float64); n = C.c_size_t()
            rc = self._L.hogb200_detect_multiscale_slot(self._h, slot, C.byref(im), r.ctypes.data,
                                                        conf.ctypes.data if confidences else None, cap, C.byref(n))
            if rc == 3:
                cap = n.value; continue
            self._ck(rc)
            return (r[:n.value].copy(), conf[:n.value].copy()) if confidences else r[:n.value].copy()

    def detectMultiScaleBatch(self, frames, nslots=1, capacity=1 << 16):
        """frames: list of images.  -> list (frame-major, slot-minor) of (N,4) int32 rect arrays."""
        keep = []
        arr = (_Image * len(frames))(*[_as_image(f, keep) for f in frames])
        while True:
            # the result buffer lives with the object: allocating (and unmapping) megabytes per call costs more than
            # the call's host work -- 0.58 ms for 16 MB against a 7.6 ms batch (tools/prof_pyoverhead.py)
            r = getattr(self, "_rect_buf", None)
            if r is None or len(r) < capacity:
                r = self._rect_buf = np.empty((capacity, 4), np.int32)
            capacity = len(r)
            offs = np.zeros(len(frames) * nslots + 1, np.uint64)
            rc = self._L.hogb200_detect_multiscale_batch(self._h, arr, len(frames), nslots, r.ctypes.data, capacity,
                                                         offs.ctypes.data)
            if rc == 3:
                capacity *= 8; continue
            self._ck(rc)
            offs = offs.astype(np.int64)
            return [r[offs[i]:offs[i + 1]].copy() for i in range(len(frames) * nslots)]

    def detectROI(self, img, locations, padding=(8, 8)):
        """cv::HOGDescriptor::detectROI -> (found (M,2) int32, confidences (len(locations),) float64)"""
        keep = []; im = _as_image(img, keep)
        loc = np.ascontiguousarray(np.asarray(locations, np.int32).reshape(-1, 2))
        found = np.zeros((max(len(loc), 1), 2), np.int32); conf = np.zeros(max(len(loc), 1), np.float64); n = C.c_size_t()
        self._ck(self._L.hogb200_detect_roi(self._h, C.byref(im), loc.ctypes.data, len(loc), int(padding[0]), int(padding[1]),
                                            found.ctypes.data, conf.ctypes.data, len(found), C.byref(n)))
        return found[:n.value].copy(), conf[:len(loc)].copy()

    def detectMultiScaleROI(self, img, rois, group_threshold=0):
        """cv::HOGDescriptor::detectMultiScaleROI.  rois: list of (scale, (K,2) locations in that level's pixels)
        -> (rects (M,4) int32, [confidences (K,) per roi])"""
        keep = []; im = _as_image(img, keep)
        locs = [np.ascontiguousarray(np.asarray(r[1], np.int32).reshape(-1, 2)) for r in rois]
        confs = [np.zeros(max(len(l), 1), np.float64) for l in locs]
        arr = (_Roi * max(len(rois), 1))(*[_Roi(float(r[0]), l.ctypes.data, len(l), c.ctypes.data) for r, l, c in zip(rois, locs, confs)])
        cap = max(sum(len(l) for l in locs), 1)
        out = np.zeros((cap, 4), np.int32); n = C.c_size_t()
        self._ck(self._L.hogb200_detect_multiscale_roi(self._h, C.byref(im), arr, len(rois), int(group_threshold), out.ctypes.data,
                                                       cap, C.byref(n)))
        return out[:n.value].copy(), [c[:len(l)].copy() for c, l in zip(confs, locs)]

    def submitBatch(self, frames, nslots=1):
        """hogb200_submit_batch: queue the batch, return a ticket for collectBatch (at most two in flight)."""
        keep = []
        arr = (_Image * len(frames))(*[_as_image(f, keep) for f in frames])
        t = C.c_int(-1)
        self._ck(self._L.hogb200_submit_batch(self._h, arr, len(frames), nslots, C.byref(t)))
        if not hasattr(self, "_inflight"):
            self._inflight = {}
        self._inflight[t.value] = (keep, arr, len(frames) * nslots)      # the frames stay alive until the batch is collected
        return t.value

    def collectBatch(self, ticket, capacity=1 << 16):
        """hogb200_collect_batch -> list (frame-major, slot-minor) of (N,4) int32 rect arrays."""
        if ticket not in getattr(self, "_inflight", {}):
            raise HogError(1, "no batch in flight under this ticket")
        keep, arr, nseg = self._inflight[ticket]
        while True:
            r = getattr(self, "_rect_buf", None)
            if r is None or len(r) < capacity:
                r = self._rect_buf = np.empty((capacity, 4), np.int32)
            offs = np.zeros(nseg + 1, np.uint64)
            rc = self._L.hogb200_collect_batch(self._h, ticket, r.ctypes.data, capacity, offs.ctypes.data)
            if rc == 3:                   # the results are kept by the library: collect again with room for them
                capacity = max(int(offs[-1]), 8 * capacity); continue
            break
        del self._inflight[ticket]
        self._ck(rc)
        offs = offs.astype(np.int64)
        return [r[offs[i]:offs[i + 1]].copy() for i in range(nseg)]

    def compute(self, img):
        """-> descriptors (windows, descriptor_size) float32"""
        keep = []; im = _as_image(img, keep)
        ds = self.getDescriptorSize()
        sx, sy = self.getWinStride()
        nwx = max((im.width - self._win[0] + sx) // sx, 0) if im.width >= self._win[0] else 0
        nwy = max((im.height - self._win[1] + sy) // sy, 0) if im.height >= self._win[1] else 0
        out = np.zeros((nwx * nwy, ds), np.float32); n = C.c_size_t()
        self._ck(self._L.hogb200_compute(self._h, C.byref(im), out.ctypes.data, out.size, 0, C.byref(n)))
        assert n.value == nwx * nwy
        return out

    # ---- stage entry points (parity tests)
    def levels(self, width, height):
        cap = 256
        sc = np.zeros(cap, np.float64); lw = np.zeros(cap, np.int32); lh = np.zeros(cap, np.int32); n = C.c_size_t()
        self._ck(self._L.hogb200_levels(self._h, width, height, sc.ctypes.data, lw.ctypes.data, lh.ctypes.data, cap, C.byref(n)))
        return sc[:n.value].copy(), lw[:n.value].copy(), lh[:n.value].copy()

    def resize(self, img, dw, dh):
        keep = []; im = _as_image(img, keep)
        out = np.zeros((dh, dw, im.channels), np.uint8)
        self._ck(self._L.hogb200_resize(self._h, C.byref(im), dw, dh, out.ctypes.data))
        return out

    def blockHistograms(self, img):
        keep = []; im = _as_image(img, keep)
        nbx = (im.width - 16) // 8 + 1 if im.width >= 16 else 0
        nby = (im.height - 16) // 8 + 1 if im.height >= 16 else 0
        out = np.zeros((max(nby, 0), max(nbx, 0), 36), np.float32); bx, by = C.c_int(), C.c_int()
        self._ck(self._L.hogb200_block_histograms(self._h, C.byref(im), out.ctypes.data, out.size, C.byref(bx), C.byref(by)))
        assert (bx.value, by.value) == (out.shape[1], out.shape[0]) or out.size == 0
        return out

    def blockHistogramsNoAlloc(self, img, out):
        """hogb200_block_histograms_noalloc: `img` a CUDA uint8 tensor, `out` a CUDA float32 tensor (blocks_y, blocks_x, 36)
        that receives the histograms; queued on the handle's stream, no allocation, no synchronisation."""
        keep = []; im = _as_image(img, keep)
        bx, by = C.c_int(), C.c_int()
        self._ck(self._L.hogb200_block_histograms_noalloc(self._h, C.byref(im), out.data_ptr(), out.numel(), C.byref(bx), C.byref(by)))
        return bx.value, by.value

    def windowScores(self, img, slot=0):
        keep = []; im = _as_image(img, keep)
        sx, sy = self.getWinStride()
        nwx = (im.width - self._win[0] + sx) // sx if im.width >= self._win[0] else 0
        nwy = (im.height - self._win[1] + sy) // sy if im.height >= self._win[1] else 0
        out = np.zeros((max(nwy, 0), max(nwx, 0)), np.float32); a, b = C.c_int(), C.c_int()
        self._ck(self._L.hogb200_window_scores(self._h, slot, C.byref(im), out.ctypes.data, out.size, C.byref(a), C.byref(b)))
        return out

    @staticmethod
    def enumerateLevels(width, height, win_size=(64, 128), scale0=1.05, nlevels=64):
        L = load_library()
        cap = 1024
        sc = np.zeros(cap, np.float64); lw = np.zeros(cap, np.int32); lh = np.zeros(cap, np.int32); n = C.c_size_t()
        rc = L.hogb200_enumerate_levels(width, height, win_size[0], win_size[1], scale0, nlevels, sc.ctypes.data,
                                        lw.ctypes.data, lh.ctypes.data, cap, C.byref(n))
        if rc:
            raise HogError(rc, L.hogb200_last_error().decode())
        return sc[:n.value].copy(), lw[:n.value].copy(), lh[:n.value].copy()

    @staticmethod
    def groupRectangles(rects, group_threshold, eps=0.2):
        L = load_library()
        r = np.ascontiguousarray(np.asarray(rects, np.int32).reshape(-1, 4)).copy()
        n = C.c_size_t()
        rc = L.hogb200_group_rectangles(r.ctypes.data, len(r), group_threshold, eps, C.byref(n))
        if rc:
            raise HogError(rc, L.hogb200_last_error().decode())
        return r[:n.value].copy()

    @staticmethod
    def groupRectanglesMeanshift(rects, weights, scales, detect_threshold=0.0, win_size=(64, 128)):
        """cv::groupRectangles_meanshift -> (rects (M,4) int32, weights (M,) float64).  Host only."""
        L = load_library()
        r = np.ascontiguousarray(np.asarray(rects, np.int32).reshape(-1, 4)).copy()
        w = np.ascontiguousarray(np.asarray(weights, np.float64).reshape(-1)).copy()
        sc = np.ascontiguousarray(np.asarray(scales, np.float64).reshape(-1))
        assert len(r) == len(w) == len(sc)
        n = C.c_size_t()
        rc = L.hogb200_group_rectangles_meanshift(r.ctypes.data, w.ctypes.data, sc.ctypes.data, len(r), detect_threshold,
                                                  win_size[0], win_size[1], C.byref(n))
        if rc:
            raise HogError(rc, "groupRectangles_meanshift: bad arguments")
        return r[:n.value].copy(), w[:n.value].copy()

    def batchGroup(self, width, height, channels=4):
        """frames per kernel launch the batch path uses for this frame shape"""
        return int(self._L.hogb200_batch_group(self._h, int(width), int(height), int(channels)))

    def groupRectanglesDevice(self, rects, group_threshold, eps=0.2):
        """cv::groupRectangles computed on the device (csrc/group.cu): same output as groupRectangles."""
        r = np.ascontiguousarray(np.asarray(rects, np.int32).reshape(-1, 4)).copy()
        n = C.c_size_t()
        self._ck(self._L.hogb200_group_rectangles_device(self._h, r.ctypes.data, len(r), group_threshold, eps, C.byref(n)))
        return r[:n.value].copy()
