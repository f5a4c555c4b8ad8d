"""The workload run under compute-sanitizer (tools/sanitize.sh): smoke()'s single small frame, an 11-frame KITTI-shaped
two-detector batch (ragged frame groups, many warps outside small levels) and an 8UC1 VGA batch -- every kernel of the
library, every staging scheme of the resize kernel, both template instances of the histogram kernel."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
from _oracle import synthetic_frame  # noqa: E402
from opencv_b200 import HOG  # noqa: E402

small = len(sys.argv) > 1 and sys.argv[1] == "small"      # racecheck is ~100x slower: fewer, smaller frames
hog = HOG.create()
det = hog.getDefaultPeopleDetector()
hog.setSVMDetector(det)
img = synthetic_frame(1, 320, 256, 4, silhouettes=2)
hog.resize(img, 305, 244); hog.resize(img, 160, 128); hog.resize(img, 100, 80)
hog.blockHistograms(img)
hog.windowScores(img)
hog.setHitThreshold(-2.0)
print("smoke-sized frame:", len(hog.detectMultiScale(img)), "rects")
veh = (np.random.RandomState(11).randn(3781) * 0.05).astype(np.float32); veh[3780] = -0.5
hog.setSVMDetector(veh, slot=1)
hog.setHitThreshold(-0.75)
w, h, n = (621, 187, 5) if small else (1242, 375, 11)
frames = [synthetic_frame(7000 + i, w, h, 4, silhouettes=4) for i in range(n)]
out = hog.detectMultiScaleBatch(frames, nslots=2)
print("kitti-shaped batch:", sum(len(o) for o in out), "rects from", n, "frames x 2 detectors")
w, h, n = (320, 240, 9) if small else (640, 480, 17)
gray = [synthetic_frame(8000 + i, w, h, 1, silhouettes=3) for i in range(n)]
out = hog.detectMultiScaleBatch(gray)
print("8UC1 batch:", sum(len(o) for o in out), "rects from", n, "frames")
bgr = [np.ascontiguousarray(f[:, :, :3]) for f in frames[:3]]
out = hog.detectMultiScaleBatch(bgr, nslots=2)
print("8UC3 batch:", sum(len(o) for o in out), "rects")
hog.setWinStride((64, 128)); d = hog.compute(img); hog.setWinStride((8, 8))
print("compute:", d.shape)
r = np.stack([np.arange(4000) % 97 * 8, np.arange(4000) % 53 * 8, np.full(4000, 64), np.full(4000, 128)], 1).astype(np.int32)
print("device groupRectangles:", len(hog.groupRectanglesDevice(r, 2, 0.2)))
import torch  # noqa: E402
t0 = hog.submitBatch(frames[:3], nslots=2); t1 = hog.submitBatch(gray[:4])
print("submit/collect:", sum(len(o) for o in hog.collectBatch(t1)), sum(len(o) for o in hog.collectBatch(t0)), "rects")
os.environ["HOGB200_GPU_POST_MIN"] = "1"
print("device tail:", sum(len(o) for o in hog.detectMultiScaleBatch(frames[:3], nslots=2)), "rects")
os.environ.pop("HOGB200_GPU_POST_MIN")
loc = np.stack([np.arange(-1, 30) * 8, np.arange(-1, 30) % 7 * 8], 1).astype(np.int32)
print("detectROI:", len(hog.detectROI(img, loc)[0]), "hits;", len(hog.detectMultiScaleROI(img, [(1.0, loc), (1.3, loc)], 1)[0]), "grouped")
d_img = torch.from_numpy(img).cuda(); out = torch.zeros((31, 39, 36), dtype=torch.float32, device="cuda")
hog.blockHistogramsNoAlloc(d_img, out); torch.cuda.synchronize()
print("no-alloc histograms:", float(out.abs().sum()) > 0)
wide = HOG.create(win_size=(128, 64)); wide.setSVMDetector((np.random.RandomState(2).randn(15 * 7 * 36 + 1) * 0.05).astype(np.float32))
print("128x64 window:", wide.windowScores(frames[0]).shape)
print("sanitize_case ok")
