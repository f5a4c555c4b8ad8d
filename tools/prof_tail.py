"""A batch of frames at a hit threshold that yields thousands of candidates per frame, through the asynchronous batch API:
the command profiled (ncu launch list) to see what the device tail -- reference order, rect scaling, groupRectangles
per (frame, detector) -- costs beside the three pipeline kernels.
usage: python tools/prof_tail.py W H channels nframes [reps]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
from _oracle import synthetic_frame  # noqa: E402
from opencv_b200 import HOG  # noqa: E402

W, H, CN, NF = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 3
hog = HOG.create(); hog.setSVMDetector(hog.getDefaultPeopleDetector())
base = [synthetic_frame(1000 + i, W, H, CN, silhouettes=6) for i in range(4)]
dev = [torch.from_numpy(np.roll(base[i % 4], 7 * (i // 4), axis=0).copy()).cuda() for i in range(NF)]
hog.setHitThreshold(-100.0); hog.setGroupThreshold(0)
_, sc = hog.detectMultiScale(dev[0], confidences=True)
target = min(4000, max(1, len(sc) // 8))
thr = float(np.sort(sc)[-target])
for hit, name in ((5.0, "no hits"), (thr, "with hits (%.2f)" % thr)):
    hog.setHitThreshold(hit); hog.setGroupThreshold(2)
    pending = hog.submitBatch(dev)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for s in range(reps):
        nxt = hog.submitBatch(dev)
        out = hog.collectBatch(pending, capacity=1 << 20)
        pending = nxt
    out = hog.collectBatch(pending, capacity=1 << 20)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    print("%-20s %8.1f frames/s  rects %d  candidates %d" % (name, NF * reps / (t1 - t0), sum(len(o) for o in out), hog.lastTimings()["candidates"]))
