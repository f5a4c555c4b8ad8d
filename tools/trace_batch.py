import os, sys
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import torch
from _oracle import synthetic_frame
from opencv_b200 import HOG
hog = HOG.create(); hog.setSVMDetector(hog.getDefaultPeopleDetector())
host = [torch.from_numpy(synthetic_frame(10 + i, 1920, 1080, 4, silhouettes=6)).pin_memory() for i in range(32)]
host_np = [h.numpy() for h in host]
for i in range(3):
    print("---- batch", i, file=sys.stderr)
    hog.detectMultiScaleBatch(host_np)
