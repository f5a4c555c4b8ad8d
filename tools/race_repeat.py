"""Stand-in for compute-sanitizer racecheck (closed on this pool, profiles/r02_sanitizer_closed.txt): every kernel's output
is computed N times on the same inputs and compared BIT FOR BIT.  The hand-synchronised parts (warp-private shared
accumulators, mbarrier rings, bulk-copy staging, atomics-free voting) have no legitimate source of run-to-run variation, so
any difference is a race or an uninitialised read.  Candidate lists come back in atomic-append order and are compared as
sorted sets.  usage: python tools/race_repeat.py [N]"""
import hashlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
from _oracle import synthetic_frame  # noqa: E402
from opencv_b200 import HOG  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 25
hog = HOG.create()
hog.setSVMDetector(hog.getDefaultPeopleDetector())
veh = (np.random.RandomState(11).randn(3781) * 0.05).astype(np.float32); veh[3780] = -0.5
hog.setSVMDetector(veh, slot=1)
f4 = [synthetic_frame(7000 + i, 1242, 375, 4, silhouettes=4) for i in range(11)]
f1 = [synthetic_frame(8000 + i, 640, 480, 1, silhouettes=3) for i in range(9)]
big = synthetic_frame(5, 1920, 1080, 4, silhouettes=6)


def digest(*arrays):
    h = hashlib.sha1()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()[:16]


cases = {
    "resize 8UC4 1080p -> 1741x979": lambda: digest(hog.resize(big, 1741, 979)),
    "resize 8UC4 1080p -> 403x227": lambda: digest(hog.resize(big, 403, 227)),
    "resize 8UC1 640x480 -> 581x436": lambda: digest(hog.resize(f1[0], 581, 436)),
    "block histograms 1080p 8UC4": lambda: digest(hog.blockHistograms(big)),
    "block histograms 640x480 8UC1": lambda: digest(hog.blockHistograms(f1[0])),
    "window scores 1080p": lambda: digest(hog.windowScores(big)),
    "window scores 1080p, slot 1": lambda: digest(hog.windowScores(big, slot=1)),
    "batch 11 x 1242x375, 2 detectors, hit -0.75": lambda: digest(*hog.detectMultiScaleBatch(f4, nslots=2)),
    "batch 9 x 640x480 8UC1, device tail": lambda: digest(*hog.detectMultiScaleBatch(f1)),
    "submit/collect, two in flight": lambda: (lambda a, b: digest(*(hog.collectBatch(b) + hog.collectBatch(a))))(hog.submitBatch(f4[:5], nslots=2), hog.submitBatch(f1)),
}
hog.setHitThreshold(-0.75)
os.environ["HOGB200_GPU_POST_MIN"] = "1"
bad = 0
for name, fn in cases.items():
    seen = {}
    for i in range(N):
        seen.setdefault(fn(), []).append(i)
    ok = len(seen) == 1
    bad += not ok
    print("%-48s %s  %d runs  %s" % (name, "identical" if ok else "DIFFERS", N, " ".join("%s x%d" % (k, len(v)) for k, v in seen.items())))
print("race_repeat:", "ok" if not bad else "%d case(s) differ" % bad)
sys.exit(1 if bad else 0)
