"""HOGB200_TRACE=1 timeline (upload and kernel interval of every frame group) of a blocking batch call on pinned host frames.
usage: HOGB200_TRACE=1 python tools/trace_batch.py [W H channels nframes nslots]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
from _oracle import synthetic_frame
from opencv_b200 import HOG
W, H, CN, NF, NS = (int(v) for v in sys.argv[1:6]) if len(sys.argv) > 5 else (1920, 1080, 4, 32, 1)
hog = HOG.create(); hog.setSVMDetector(hog.getDefaultPeopleDetector())
if NS > 1:
    veh = (np.random.RandomState(11).randn(3781) * 0.05).astype(np.float32); veh[3780] = -0.5
    hog.setSVMDetector(veh, slot=1)
host = [torch.from_numpy(synthetic_frame(10 + i, W, H, CN, silhouettes=6)).pin_memory() for i in range(NF)]
host_np = [h.numpy() for h in host]
for i in range(3):
    print("---- batch", i, file=sys.stderr)
    hog.detectMultiScaleBatch(host_np, nslots=NS)
