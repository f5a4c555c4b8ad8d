"""Stage timers of the batch path with frames in pinned host memory vs resident in HBM (last group of the batch)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
from _oracle import synthetic_frame  # noqa: E402
from opencv_b200 import HOG  # noqa: E402

hog = HOG.create()
hog.setSVMDetector(hog.getDefaultPeopleDetector())
host = [torch.from_numpy(synthetic_frame(10 + i, 1920, 1080, 4, silhouettes=6)).pin_memory() for i in range(32)]
dev = [h.cuda() for h in host]
host_np = [h.numpy() for h in host]
for name, frames in (("device", dev), ("host", host_np)):
    for timing in (False, True):
        hog.enableTiming(timing)
        for i in range(3):
            hog.detectMultiScaleBatch(frames)
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(5):
            hog.detectMultiScaleBatch(frames)
        e1.record(); torch.cuda.synchronize()
        t = hog.lastTimings()
        print(name, "timing" if timing else "plain", "ms/batch %.3f" % (e0.elapsed_time(e1) / 5),
              {k: round(v, 4) for k, v in t.items() if k.endswith("_ms")})
