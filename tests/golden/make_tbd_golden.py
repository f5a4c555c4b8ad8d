"""Records tests/golden/tbd_golden.npz from the UNMODIFIED reference tracker (oracle/_ref/libcvref.so, built by
oracle/Makefile from /root/reference): the scalar state of every track after every step of one seeded sequence.
Run in the build container:  python tests/golden/make_tbd_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_tbd import RefTracker, scene  # noqa: E402
from opencv_b200 import tbd  # noqa: E402

seed, nframes, nobj = 1234, 48, 12
args = (10.0, 16, 4, 0.3, 0.2, 1)
ref = RefTracker(tbd.TbdArgs(*args))
out = dict(seed=seed, nframes=nframes, nobj=nobj, args=np.array(args, np.float64))
for k, (f, dets) in enumerate(scene(seed, nframes, nobj, True)):
    ref.step(dets, f)
    st = ref.state()
    out["ints_%d" % k] = np.array([s[0] for s in st], np.int32).reshape(-1, 13)
    out["dbl_%d" % k] = np.array([s[1] for s in st], np.float64).reshape(-1, 3)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "tbd_golden.npz"), **out)
print("tracks at the end:", len(st))
