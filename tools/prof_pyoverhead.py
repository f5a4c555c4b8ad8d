"""Host-side cost of one detectMultiScaleBatch call around the C entry point (ctypes descriptors, result split)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from _oracle import synthetic_frame
from opencv_b200 import HOG
from opencv_b200.hog import _as_image, _Image
hog = HOG.create(); hog.setSVMDetector(hog.getDefaultPeopleDetector())
dev = [torch.from_numpy(synthetic_frame(10 + i % 4, 1920, 1080, 4)).cuda() for i in range(32)]
for _ in range(3): hog.detectMultiScaleBatch(dev)
t = []
for _ in range(10):
    t0 = time.perf_counter(); keep = []; arr = (_Image * 32)(*[_as_image(f, keep) for f in dev]); t1 = time.perf_counter()
    r = np.zeros((1 << 20, 4), np.int32); offs = np.zeros(33, np.uint64); t2 = time.perf_counter()
    hog._L.hogb200_detect_multiscale_batch(hog._h, arr, 32, 1, r.ctypes.data, 1 << 20, offs.ctypes.data); t3 = time.perf_counter()
    o = offs.astype(np.int64); out = [r[o[i]:o[i + 1]].copy() for i in range(32)]; t4 = time.perf_counter()
    t.append((t1 - t0, t2 - t1, t3 - t2, t4 - t3))
print("ms: descriptors %.3f  alloc %.3f  C call %.3f  split %.3f" % tuple(1e3 * np.median(np.array(t), 0)))
