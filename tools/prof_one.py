"""One 1080p BGRA frame through detectMultiScale a few times: the command profiled by ncu (see profiles/README)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
from _oracle import synthetic_frame  # noqa: E402
from opencv_b200 import HOG  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4
W, H = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1920, 1080)
hog = HOG.create()
hog.setSVMDetector(hog.getDefaultPeopleDetector())
frames = [torch.from_numpy(synthetic_frame(10 + i, W, H, 4, silhouettes=6)).cuda() for i in range(2)]
hog.enableTiming(True)
for i in range(n):
    r = hog.detectMultiScale(frames[i % 2])
    t = hog.lastTimings()
print("rects", len(r), {k: round(v, 4) if isinstance(v, float) else v for k, v in t.items()})
