"""Host-side timeline of the asynchronous batch API: wall-clock of every submit / collect call of a pipelined loop
(two batches in flight) on pinned host frames, with a hit threshold that yields thousands of candidates per frame.
usage: python tools/prof_async.py W H frames_per_batch [steps]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
from _oracle import synthetic_frame  # noqa: E402
from opencv_b200 import HOG  # noqa: E402

W, H, NB = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 6
hog = HOG.create(); hog.setSVMDetector(hog.getDefaultPeopleDetector())
base = [synthetic_frame(1000 + i, W, H, 4, silhouettes=6) for i in range(4)]
pinned = [torch.from_numpy(np.roll(base[i % 4], 7 * (i // 4), axis=0).copy()).pin_memory() for i in range(NB)]
frames = [p.numpy() for p in pinned]
hog.setHitThreshold(-100.0); hog.setGroupThreshold(0)
_, sc = hog.detectMultiScale(pinned[0].cuda(), confidences=True)
thr = float(np.sort(sc)[-4000])
for hit, name in ((thr, "with hits (%.2f)" % thr), (5.0, "no hits")):
    hog.setHitThreshold(hit); hog.setGroupThreshold(2)
    print("----", name)
    t0 = time.perf_counter()
    pending = hog.submitBatch(frames)
    print("submit %.1f ms" % (1e3 * (time.perf_counter() - t0)))
    for s in range(steps):
        t0 = time.perf_counter(); nxt = hog.submitBatch(frames); t1 = time.perf_counter()
        out = hog.collectBatch(pending, capacity=1 << 20); t2 = time.perf_counter()
        print("step %d: submit %.1f ms  collect %.1f ms  rects %d  candidates %d" % (s, 1e3 * (t1 - t0), 1e3 * (t2 - t1), sum(len(o) for o in out), hog.lastTimings()["candidates"]))
        pending = nxt
    hog.collectBatch(pending, capacity=1 << 20)
