"""The other BASELINE.json configs through the batch path, with stage timings (not bench lines):
KITTI 1242x375 with two detectors, 3840x2160, and the dense stress (1080p, scale 1.02, 128 levels, hit 0, group 0)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
from _oracle import synthetic_frame  # noqa: E402
from opencv_b200 import HOG  # noqa: E402


def run(name, W, H, nframes, nslots=1, scale=1.05, nlevels=64, hit=0.0, group=2, reps=4):
    hog = HOG.create()
    det = hog.getDefaultPeopleDetector()
    hog.setSVMDetector(det, 0)
    if nslots > 1:
        rng = np.random.RandomState(3)
        hog.setSVMDetector((det + 0.02 * rng.standard_normal(det.shape)).astype(np.float32), 1)   # stand-in vehicle model
    hog.setScaleFactor(scale); hog.setNumLevels(nlevels); hog.setHitThreshold(hit); hog.setGroupThreshold(group)
    frames = [torch.from_numpy(synthetic_frame(10 + i, W, H, 4, silhouettes=4)).cuda() for i in range(nframes)]
    hog.enableTiming(True)
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(reps):
        if i == reps - 1: ev0.record()
        r = hog.detectMultiScaleBatch(frames, nslots=nslots, capacity=1 << 22)
    ev1.record(); torch.cuda.synchronize()
    t = hog.lastTimings()
    ms = ev0.elapsed_time(ev1)
    print("%-28s %dx%d x%d: %.3f ms/frame (%.0f fps)  last group: resize %.3f hist %.3f svm %.3f ms, %d levels px %d windows %d cand %d rects %d" % (
        name, W, H, nframes, ms / nframes, 1000 * nframes / ms, t["resize_ms"], t["hist_ms"], t["svm_ms"],
        len(hog.levels(W, H)), t["pyramid_pixels"], t["windows"], t["candidates"], sum(len(x) for x in r)))


run("kitti 2 detectors", 1242, 375, 32, nslots=2)
run("4k", 3840, 2160, 8)
run("dense 1.02/128 hit0 group0", 1920, 1080, 8, scale=1.02, nlevels=128, hit=0.0, group=0)
run("1080p default", 1920, 1080, 32)
