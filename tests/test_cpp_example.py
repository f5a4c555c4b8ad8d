"""include/hogb200.hpp (header-only C++ mirror of cv::cuda::HOG / cv::tbd::Tracker over the C ABI) and
examples/tbd_loop.cpp (example_gpu_tbd's frame loop): compiles with plain g++, links against the product library; without
a GPU it must fail loudly, on a GPU it must detect and track the synthetic walkers."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "examples", "tbd_loop")


def build_example():
    lib = os.path.join(ROOT, "opencv_b200")
    subprocess.check_call(["g++", "-std=c++14", "-O2", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "examples", "tbd_loop.cpp"), "-o", EXE, "-L" + lib, "-lhogb200",
                           "-Wl,-rpath," + lib])


def test_cpp_mirror_compiles_and_fails_loudly_without_gpu():
    import torch
    build_example()
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    p = subprocess.run([EXE], capture_output=True, text=True)
    assert p.returncode == 1 and "no CPU fallback" in p.stderr


@pytest.mark.gpu
def test_cpp_example_tracks_the_walkers():
    build_example()
    p = subprocess.run([EXE], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr
    lines = p.stdout.splitlines()
    assert sum(l.startswith("frame") for l in lines) == 12
    tracks = [l for l in lines if l.startswith("track")]
    ages = [int(l.split("age ")[1].split(",")[0]) for l in tracks]
    assert len(tracks) >= 1 and max(ages) >= 6, p.stdout            # a walker is followed through most of the 12 frames
    # the same loop on the asynchronous batch API (two batches in flight): identical detections and tracks
    q = subprocess.run([EXE, "--pipelined"], capture_output=True, text=True, timeout=120)
    assert q.returncode == 0, q.stderr
    assert q.stdout == p.stdout
