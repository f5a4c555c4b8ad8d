import os, sys, threading, time
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import torch
from _oracle import synthetic_frame
from opencv_b200 import HOG
hog = HOG.create(); hog.setSVMDetector(hog.getDefaultPeopleDetector())
host = [torch.from_numpy(synthetic_frame(10 + i, 1920, 1080, 4, silhouettes=6)).pin_memory() for i in range(32)]
dev = [h.cuda() for h in host]
side = torch.cuda.Stream()
big = torch.empty(33177600, dtype=torch.uint8).pin_memory(); dbig = torch.empty(33177600, dtype=torch.uint8, device='cuda')
def run(bg):
    for i in range(3): hog.detectMultiScaleBatch(dev)
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    if bg:
        with torch.cuda.stream(side):
            for _ in range(120): dbig.copy_(big, non_blocking=True)     # ~72 ms of back-to-back H2D
    e0.record()
    for i in range(5): hog.detectMultiScaleBatch(dev)
    e1.record(); torch.cuda.synchronize()
    print("background H2D" if bg else "quiet", "ms/batch %.3f" % (e0.elapsed_time(e1)/5))
run(False); run(True); run(False); run(True)
