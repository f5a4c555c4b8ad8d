"""Where the cycles of svm_kernel's consumer warps go (experiments only).  Needs a library built with
-DHOGB200_SVM_TIMING (tools/build_variant.sh svmtime -DHOGB200_SVM_TIMING), selected with HOGB200_LIB; run with
HOGB200_LANES=1 so that two svm launches never add to the counters at once."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
from _oracle import synthetic_frame  # noqa: E402
from opencv_b200 import HOG  # noqa: E402
from opencv_b200 import hog as hogmod  # noqa: E402

W, H = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1920, 1080)
group = int(os.environ.get("HOGB200_GROUP", "4"))
hog = HOG.create()
hog.setSVMDetector(hog.getDefaultPeopleDetector())
frames = [torch.from_numpy(synthetic_frame(10 + i, W, H, 4, silhouettes=6)).cuda() for i in range(group)]
lib = ctypes.CDLL(hogmod.lib_path())      # same path -> the handle the wrapper already loaded
fn = lib.hogb200_debug_svm_timing
fn.argtypes = [ctypes.c_void_p, ctypes.c_int]
for i in range(3):
    hog.detectMultiScaleBatch(frames)
torch.cuda.synchronize()
assert fn(None, 1) == 0
reps = 5
hog.enableTiming(True)
for i in range(reps):
    hog.detectMultiScaleBatch(frames)
    t = hog.lastTimings()
torch.cuda.synchronize()
out = np.zeros((160, 8, 8), dtype=np.uint64)
assert fn(out.ctypes.data, 0) == 0
a = out[:148].astype(np.float64) / reps
names = ["dead tiles", "first wait", "step waits", "steps", "epilogue", "rest wait", "live tiles", "total"]
tot = a[:, :, 7]
print("timings", {k: round(v, 4) if isinstance(v, float) else v for k, v in t.items()})
print("kernel cycles per warp: mean %.0f  min %.0f  max %.0f  (max CTA mean %.0f, min CTA mean %.0f)" %
      (tot.mean(), tot.min(), tot.max(), tot.mean(1).max(), tot.mean(1).min()))
for i in (0, 1, 2, 3, 4, 5):
    print("%-12s %6.1f %% of the longest warp's cycles (mean over warps); by warp: %s" %
          (names[i], 100 * a[:, :, i].mean() / tot.max(), " ".join("%5.1f" % (100 * a[:, w, i].mean() / tot.max()) for w in range(8))))
idle = tot.max() - tot
print("%-12s %6.1f %%  (warp finished before the longest one)" % ("tail", 100 * idle.mean() / tot.max()))
lt = a[:, :, 6]
print("live tiles per warp: mean %.2f; steps cycles per live tile-step: %.0f (FFMA2 pipe alone: 1008 per warp, 2016 for two warps sharing a sub-partition)" %
      (lt.mean(), a[:, :, 3].sum() / max(1.0, lt.sum()) / 16))
print("by CTA quartile of total:", np.percentile(tot.mean(1), [0, 25, 50, 75, 100]).round(0))
