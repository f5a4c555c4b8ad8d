"""groupRectangles: host restatement vs device path (csrc/group.cu) on synthetic candidate clouds; wall-clock ms."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
from test_gpu_parity import _rect_cloud  # noqa: E402
from opencv_b200 import HOG  # noqa: E402

hog = HOG.create()
hog.groupRectanglesDevice(_rect_cloud(1, 100, 4), 2)
for n, k in ((2048, 30), (20000, 150), (100000, 800), (261558, 2000)):
    r = _rect_cloud(n, n, k)
    t0 = time.perf_counter(); a = HOG.groupRectangles(r, 2, 0.2); t1 = time.perf_counter()
    b = hog.groupRectanglesDevice(r, 2, 0.2); t2 = time.perf_counter()
    b = hog.groupRectanglesDevice(r, 2, 0.2); t3 = time.perf_counter()
    print("n=%7d  groups %5d  host %9.2f ms   device %8.2f ms (second call %8.2f ms)  equal=%s" % (
        n, len(a), (t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3, np.array_equal(a, b)))
