"""ctypes mirror of cv::tbd (modules/trackingbydetection/include/opencv2/tbd.hpp:25-184) over the C ABI of
include/tbdb200.h: TbdArgs, Detection, Tracker.performTrackingStep / getTracks / setTracks / reset and the per-frame
metric vectors, with the reference's names.  Host code; lives in the same library as the detector."""
import ctypes as C

import numpy as np

from .hog import HogError, load_library


class TbdArgs(C.Structure):
    """cv::tbd::TbdArgs (tbd.hpp:25-41)"""
    _fields_ = [("costOfNonAssignment", C.c_double), ("timeWindowSize", C.c_uint), ("trackAgeThreshold", C.c_uint),
                ("trackVisibilityThreshold", C.c_double), ("trackConfidenceThreshold", C.c_double),
                ("shouldStoreMetrics", C.c_int)]


class Detection(C.Structure):
    """cv::tbd::Detection (tbd.hpp:79-93); Detection(id, frame_id, (x, y, w, h), confidence)"""
    _fields_ = [("id", C.c_int), ("frame_id", C.c_int), ("x", C.c_int), ("y", C.c_int), ("width", C.c_int),
                ("height", C.c_int), ("confidence", C.c_double)]

    def __init__(self, id=-1, frame_id=0, bbox=(0, 0, 0, 0), confidence=-1.0):
        super().__init__(int(id), int(frame_id), int(bbox[0]), int(bbox[1]), int(bbox[2]), int(bbox[3]), float(confidence))


class Track(C.Structure):
    """scalar state of a cv::tbd::Track (tbd.hpp:99-126)"""
    _fields_ = [("id", C.c_uint), ("age", C.c_uint), ("totalVisibleCount", C.c_uint), ("historyLen", C.c_int),
                ("maxConfidence", C.c_double), ("avgConfidence", C.c_double), ("bboxOverlap", C.c_double),
                ("pred_x", C.c_int), ("pred_y", C.c_int), ("pred_width", C.c_int), ("pred_height", C.c_int),
                ("x", C.c_int), ("y", C.c_int), ("width", C.c_int), ("height", C.c_int), ("lastFrame", C.c_int)]

    @property
    def predPosition(self):
        return (self.pred_x, self.pred_y, self.pred_width, self.pred_height)

    @property
    def bbox(self):
        return (self.x, self.y, self.width, self.height)


class TrajectoryFrame(C.Structure):
    """what cv::tbd::Trajectory holds for one frame (tbd.hpp:60-68)"""
    _fields_ = [("hasInfo", C.c_int), ("isTracked", C.c_int), ("trackId", C.c_int), ("trackPos", C.c_int * 4),
                ("predPos", C.c_int * 4), ("bboxOverlap", C.c_double)]


class FrameMetrics(C.Structure):
    _fields_ = [("truePositives", C.c_int), ("falseNegatives", C.c_int), ("falsePositives", C.c_int),
                ("groundTruths", C.c_int), ("numMatches", C.c_int), ("bboxOverlap", C.c_double)]


_BOUND = False


def _lib():
    global _BOUND
    L = load_library()
    if not _BOUND:
        L.tbdb200_last_error.restype = C.c_char_p
        L.tbdb200_create.argtypes = [C.POINTER(TbdArgs), C.POINTER(C.c_void_p)]
        L.tbdb200_destroy.argtypes = [C.c_void_p]; L.tbdb200_destroy.restype = None
        L.tbdb200_reset.argtypes = [C.c_void_p]
        L.tbdb200_set_bounds.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
        L.tbdb200_step.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.tbdb200_num_tracks.argtypes = [C.c_void_p]
        L.tbdb200_get_tracks.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
        L.tbdb200_track_history.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
        L.tbdb200_copy_tracks.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        L.tbdb200_set_tracks.argtypes = [C.c_void_p, C.c_void_p]
        L.tbdb200_trackset_destroy.argtypes = [C.c_void_p]; L.tbdb200_trackset_destroy.restype = None
        L.tbdb200_metrics.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
        L.tbdb200_trajectory_add_position.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.tbdb200_trajectory_get.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(TrajectoryFrame)]
        L.tbdb200_trajectory_clear.argtypes = [C.c_void_p]
        L.tbdb200_bbox_overlap.argtypes = [C.c_void_p, C.c_void_p]; L.tbdb200_bbox_overlap.restype = C.c_double
        _BOUND = True
    return L


class TrackSet:
    """A copy of a tracker's vector<Track> (what the sample keeps in its history buffers)."""

    def __init__(self, lib, handle):
        self._L, self._h = lib, handle

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.tbdb200_trackset_destroy(self._h)
            self._h = None


class Tracker:
    """cv::tbd::Tracker (tbd.hpp:131-184)"""

    def __init__(self, args):
        self._L = _lib()
        self.args = args
        h = C.c_void_p()
        self._ck(self._L.tbdb200_create(C.byref(args), C.byref(h)))
        self._h = h

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.tbdb200_destroy(self._h)
            self._h = None

    def _ck(self, rc):
        if rc != 0:
            raise HogError(rc, (self._L.tbdb200_last_error() or b"").decode())

    def reset(self):
        self._ck(self._L.tbdb200_reset(self._h))

    def setBounds(self, xmin, xmax, ymin, ymax):
        self._ck(self._L.tbdb200_set_bounds(self._h, xmin, xmax, ymin, ymax))

    def performTrackingStep(self, detections, frame_id):
        """detections: list of Detection.  Returns, per detection, the id of the track it was assigned to (-1: none,
        the detection started a new track)."""
        n = len(detections)
        arr = (Detection * max(n, 1))(*detections)
        out = np.full(max(n, 1), -1, np.int32)
        self._ck(self._L.tbdb200_step(self._h, arr, n, int(frame_id), out.ctypes.data_as(C.c_void_p)))
        return out[:n]

    def getTracks(self):
        n = self._L.tbdb200_num_tracks(self._h)
        arr = (Track * max(n, 1))()
        cnt = C.c_int()
        self._ck(self._L.tbdb200_get_tracks(self._h, arr, n, C.byref(cnt)))
        return list(arr[:n])

    def trackHistory(self, index):
        """(frames[n], bboxes[n,4], scores[n]) of track `index`"""
        cnt = C.c_int()
        self._ck(self._L.tbdb200_track_history(self._h, index, None, None, None, 0, C.byref(cnt)))
        n = cnt.value
        fr = np.zeros(n, np.int32); bx = np.zeros((n, 4), np.int32); sc = np.zeros(n, np.float64)
        self._ck(self._L.tbdb200_track_history(self._h, index, fr.ctypes.data_as(C.c_void_p), bx.ctypes.data_as(C.c_void_p),
                                               sc.ctypes.data_as(C.c_void_p), n, C.byref(cnt)))
        return fr, bx, sc

    def copyTracks(self):
        h = C.c_void_p()
        self._ck(self._L.tbdb200_copy_tracks(self._h, C.byref(h)))
        return TrackSet(self._L, h)

    def setTracks(self, trackset):
        self._ck(self._L.tbdb200_set_tracks(self._h, trackset._h if trackset is not None else None))

    def addPosition(self, id, frame, bbox):
        """Trajectory::addPosition of ground-truth object `id`"""
        self._ck(self._L.tbdb200_trajectory_add_position(self._h, int(id), int(frame), (C.c_int * 4)(*map(int, bbox))))

    def trajectoryFrame(self, id, frame):
        out = TrajectoryFrame()
        self._ck(self._L.tbdb200_trajectory_get(self._h, int(id), int(frame), C.byref(out)))
        return out

    def clearTrajectories(self):
        self._ck(self._L.tbdb200_trajectory_clear(self._h))

    def metrics(self):
        cnt = C.c_int()
        self._ck(self._L.tbdb200_metrics(self._h, None, 0, C.byref(cnt)))
        arr = (FrameMetrics * max(cnt.value, 1))()
        self._ck(self._L.tbdb200_metrics(self._h, arr, cnt.value, C.byref(cnt)))
        return list(arr[:cnt.value])


def computeBoundingBoxOverlap(a, b):
    """tbd.cpp:1083-1106"""
    aa = (C.c_int * 4)(*map(int, a)); bb = (C.c_int * 4)(*map(int, b))
    return _lib().tbdb200_bbox_overlap(aa, bb)
