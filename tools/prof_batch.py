"""A group of 1080p BGRA frames through detectMultiScaleBatch a few times: the command profiled by ncu for the
batch path (one launch of each kernel per group of frames)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
from _oracle import synthetic_frame  # noqa: E402
from opencv_b200 import HOG  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4
group = int(os.environ.get("HOGB200_GROUP", "4"))
W, H = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1920, 1080)
hog = HOG.create()
hog.setSVMDetector(hog.getDefaultPeopleDetector())
frames = [torch.from_numpy(synthetic_frame(10 + i, W, H, 4, silhouettes=6)).cuda() for i in range(group)]
hog.enableTiming(True)
for i in range(n):
    r = hog.detectMultiScaleBatch(frames)
    t = hog.lastTimings()
print("rects", sum(len(x) for x in r), {k: round(v, 4) if isinstance(v, float) else v for k, v in t.items()})
