#!/usr/bin/env python
"""bench.py -- HOG detectMultiScale frames/s at 1080p (BASELINE.json metric), one process per GPU.

A step = one pass of the hot path (pyramid -> fused gradient+histogram -> SVM+compaction -> rects + grouping) over one
batch of BATCH synthetic 1920x1080 8UC4 frames (BASELINE configs[1]: people detector, nlevels=64, scale 1.05).

  value          frames/s, whole job, frames already resident in HBM when the timed region starts
  e2e            same metric through the public call with PINNED HOST frames (H2D + result D2H inside the timed region)
  e2e_with_hits  the e2e measurement again at a hit threshold that makes every frame yield thousands of candidates and
                 grouped rects, so that compaction, the candidate D2H, rect scaling and groupRectangles are timed too
  sustained      `value` measured over seconds instead of 20 steps (clocks settle under load)
  roofline       fused gradient+histogram kernel: algorithmic bytes ((cn + 2.25) B per pyramid pixel) / its CUDA-event
                 time against the measured HBM peak, and its warp-instruction count against the issue peak
  cpu_baseline   the reference's cv::HOGDescriptor::detectMultiScale (oracle/_ref) on the host cores, bounded sample

`--impl reference` times only the reference CPU implementation (rank 0), same metric and config.
"""
import argparse
import hashlib
import json
import os
import statistics
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

W, H, CN = 1920, 1080, 4
NLEVELS, SCALE0, HIT, GROUP = 64, 1.05, 0.0, 2
NSLOTS = 1
BATCH = 32                       # frames per step per GPU: 32 x 8.3 MB = 265 MB of distinct input > 126 MB L2
WIN = (64, 128)
METRIC = "HOG detectMultiScale frames/s at 1080p"
UNIT = "frames/s"
WORKLOAD = ("cv::cuda::HOG people detector, synthetic 1920x1080 8UC4, nlevels=64, scale 1.05, "
            "hit_threshold 0, group_threshold 2 (BASELINE configs[1])")
PARALLELISM = "frame-sharded (frame g -> rank g % N), no collective on the data path, host gather of boxes"

# The driver's line is the default workload (BASELINE.json's metric is quoted on configs[1]).  north_star also asks for
# frames/s "on synthetic frames of each named shape": --workload selects the other BASELINE configs; the line keeps the
# same keys, `metric` names the shape.  (shape, channels, nlevels, scale, hit, group, detectors, frames per step per GPU)
WORKLOADS = {
    "1080p": None,
    "vga": dict(W=640, H=480, CN=1, NLEVELS=64, SCALE0=1.05, HIT=0.0, GROUP=2, NSLOTS=1, BATCH=512,
                name="cv::HOGDescriptor default people detector, synthetic 640x480 8UC1, nlevels=64, scale 1.05, hit 0, "
                     "group 2 (BASELINE configs[0])", metric="HOG detectMultiScale frames/s at 640x480 8UC1"),
    "kitti": dict(W=1242, H=375, CN=4, NLEVELS=64, SCALE0=1.05, HIT=0.0, GROUP=2, NSLOTS=2, BATCH=128,
                  name="example_gpu_tbd-style 1242x375 8UC4 sequence, pedestrian + (synthetic) vehicle detector scored in "
                       "one pass, nlevels=64, scale 1.05, hit 0, group 2 (BASELINE configs[2])",
                  metric="HOG detectMultiScale frames/s at 1242x375, two detectors"),
    "4k": dict(W=3840, H=2160, CN=4, NLEVELS=64, SCALE0=1.05, HIT=0.0, GROUP=2, NSLOTS=1, BATCH=64,
               name="synthetic 3840x2160 8UC4 frame stream, 64 frames per GPU per step (512 at 8 GPUs), nlevels=64, "
                    "scale 1.05, hit 0, group 2 (BASELINE configs[3])", metric="HOG detectMultiScale frames/s at 3840x2160"),
    "dense": dict(W=1920, H=1080, CN=4, NLEVELS=128, SCALE0=1.02, HIT=0.0, GROUP=0, NSLOTS=1, BATCH=32,
                  name="dense-pyramid stress: synthetic 1920x1080 8UC4, scale 1.02, nlevels=128, hit 0, group 0 "
                       "(BASELINE configs[4])", metric="HOG detectMultiScale frames/s at 1080p, scale 1.02 / 128 levels"),
}


def select_workload(name):
    w = WORKLOADS[name]
    if w is None:
        return
    g = globals()
    for k in ("W", "H", "CN", "NLEVELS", "SCALE0", "HIT", "GROUP", "NSLOTS", "BATCH"):
        g[k] = w[k]
    g["METRIC"] = w["metric"]
    g["WORKLOAD"] = w["name"]


def make_config():
    """The same dict from both arms (the driver compares them): the workload and its work sizes, computed from the level
    rule on the host (hogb200_enumerate_levels needs no GPU)."""
    from opencv_b200 import HOG
    sc, lw, lh = HOG.enumerateLevels(W, H, WIN, SCALE0, NLEVELS)
    px = int(sum(int(a) * int(b) for a, b in zip(lw, lh)))
    wins = int(sum(((int(a) - WIN[0]) // 8 + 1) * ((int(b) - WIN[1]) // 8 + 1) for a, b in zip(lw, lh)
                   if a >= WIN[0] and b >= WIN[1]))
    mb = BATCH * W * H * CN / 1e6
    return {"workload": WORKLOAD, "frames_per_step_per_gpu": BATCH, "levels": len(sc), "pyramid_pixels": px,
            "windows": wins, "detectors": NSLOTS, "parallelism": PARALLELISM,
            "l2": "inputs larger than L2 (%d distinct frames = %.0f MB per step per GPU)" % (BATCH, mb)}


def git_blob_hash(path):
    """`git hash-object` of a file (the ncu capture is stamped with the hash of the kernel source it was taken from)."""
    data = open(path, "rb").read()
    return hashlib.sha1(b"blob %d\0" % len(data) + data).hexdigest()


def bind_near_gpu(local_rank):
    """Pin this rank's host threads to the CPUs of its GPU's NUMA node BEFORE any pinned buffer is allocated, so
    that the frames it uploads live in memory local to the GPU's PCIe root (first touch).  With 8 ranks streaming
    8 MB frames, buffers that all sit on one socket make the upload, not the kernels, the limit.  Best effort."""
    try:
        import torch
        p = torch.cuda.get_device_properties(local_rank)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        txt = open("/sys/bus/pci/devices/%s/local_cpulist" % bdf).read().strip()
        cpus = set()
        for part in txt.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
        return "%s numa cpus %s" % (bdf, txt)
    except Exception as e:      # no sysfs entry, no permission, older torch: stay unbound
        return "unbound (%s)" % e


def make_frames(n, seed0):
    from _oracle import synthetic_frame
    base = [synthetic_frame(seed0 + i, W, H, CN, silhouettes=6) for i in range(min(n, 8))]
    frames = []
    for i in range(n):                     # distinct bytes per frame without paying n full syntheses
        f = np.roll(base[i % len(base)], shift=(7 * (i // len(base)), 13 * (i // len(base))), axis=(0, 1)).copy()
        frames.append(f)
    return frames


class ClockSampler:
    """SM clock and throttle reasons of this rank's GPU DURING the timed region: NVML polled in-process every 10 ms
    (nvidia-smi as a subprocess takes longer to start on an 8-GPU box than a timed region lasts)."""
    REASONS = (("hw_slowdown", 0x8), ("sw_power_cap", 0x4), ("sw_thermal_slowdown", 0x20), ("hw_thermal_slowdown", 0x40))

    def __init__(self, gpu, period=0.01):
        self.gpu = gpu; self.sm = []; self.mx = None; self.reasons = set(); self.stop_flag = False; self.t = None; self.err = None
        self.period = period

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            idx = self.gpu
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            if vis:                                           # NVML numbers the physical devices
                try:
                    idx = int(vis.split(",")[self.gpu])
                except ValueError:
                    pass
            h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))

            def loop():
                while not self.stop_flag:
                    try:
                        self.sm.append(float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
                        r = pynvml.nvmlDeviceGetCurrentClocksEventReasons(h) if hasattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons") \
                            else pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                        for name, bit in self.REASONS:
                            if r & bit:
                                self.reasons.add(name)
                    except Exception as e:
                        self.err = str(e)
                    time.sleep(self.period)
            self.t = threading.Thread(target=loop, daemon=True); self.t.start()
        except Exception as e:
            self.err = str(e)
        return self

    def stop(self):
        self.stop_flag = True
        if self.t:
            self.t.join(timeout=1.0)
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.mx, "reasons": ["NVML unavailable: %s" % self.err], "samples": 0}
        return {"sm_mhz": statistics.median(self.sm), "sm_max_mhz": self.mx, "reasons": sorted(self.reasons),
                "samples": len(self.sm)}


def host_cores():
    """Threads the CPU arm may use: every core this process may run on.  torchrun exports OMP_NUM_THREADS=1 to its
    workers; that is a default for torch's own pools, not a property of the box, and is deliberately overridden."""
    return max(1, len(os.sched_getaffinity(0)))


def unleash_openmp(n):
    """Make the reference's `#pragma omp parallel for num_threads(n)` (core/src/parallel.cpp:523) really get n threads:
    the reference switches OpenMP's dynamic adjustment on at load time (parallel.cpp:403-411), and libgomp then caps every
    team at the OMP_NUM_THREADS it saw at start-up -- 1 under torchrun.  So: the environment before libgomp loads, and
    dynamic adjustment off + the thread count on the runtime libcvref.so is linked against."""
    os.environ["OMP_NUM_THREADS"] = str(n)
    os.environ["OMP_DYNAMIC"] = "FALSE"
    try:
        import ctypes
        g = ctypes.CDLL("libgomp.so.1", mode=ctypes.RTLD_GLOBAL)
        g.omp_set_dynamic(0)
        g.omp_set_num_threads(int(n))
    except OSError:
        pass


def cpu_reference_fps(frames, nthreads=None):
    """The reference's own CPU path on this box's host cores (oracle/_ref), else the C port.  Test/bench
    infrastructure only -- never on the product path."""
    from _oracle import Ref, Port
    if Ref.available():
        unleash_openmp(nthreads or host_cores())
        ref = Ref()
        ref.set_num_threads(nthreads or host_cores())
        unleash_openmp(nthreads or host_cores())      # again: loading the library switched dynamic adjustment back on
        det = ref.people_detector(0)
        cores = ref.get_num_threads()
        t = []
        for f in frames:
            t0 = time.perf_counter()
            for _ in range(NSLOTS):                   # one stock detector object per SVM model
                ref.detect_multiscale_stock(f, det, hit_threshold=HIT, win_stride=(8, 8), scale0=SCALE0, nlevels=NLEVELS,
                                            final_threshold=float(GROUP))
            t.append(time.perf_counter() - t0)
        return {"kind": "reference", "cores": cores, "times": t}
    port = Port()
    from opencv_b200 import load_library
    import ctypes as C
    L = load_library(); p, n = C.c_void_p(), C.c_size_t()
    L.hogb200_default_people_detector(64, 128, C.byref(p), C.byref(n))
    det = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(n.value,)).copy()
    t = []
    for f in frames:
        t0 = time.perf_counter()
        port.detect_multiscale(f, det, hit_threshold=HIT, scale0=SCALE0, nlevels=NLEVELS, group_threshold=GROUP)
        t.append(time.perf_counter() - t0)
    return {"kind": "port", "cores": 1, "times": t}


def run_reference(args, rank):
    if rank != 0:
        return
    frames = make_frames(2, 1000)
    r = cpu_reference_fps([frames[i % 2] for i in range(args.warmup)])          # warm-up
    r = cpu_reference_fps([frames[i % 2] for i in range(args.steps)])
    total = sum(r["times"])
    fps = args.steps / total
    line = {"impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": make_config(),
            "sample_frames_per_step": 1,
            "note": "stock cv::HOGDescriptor::detectMultiScale (CPU, BGR view of the same BGRA frames); each step is a bounded "
                    "sample of the workload: 1 frame",
            "cpu_baseline": {"value": fps, "unit": UNIT, "cores": r["cores"], "host_cores": host_cores(), "kind": r["kind"],
                             "sample": "%d x 1 frame %dx%d, %s threads of %d host cores (OpenMP parallel_for_ over levels)"
                                       % (args.steps, W, H, r["cores"], host_cores())},
            "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--quick", action="store_true", help="skip the with-hits, pass-all and sustained legs (profiling runs)")
    ap.add_argument("--sustain", type=float, default=2.0, help="seconds of the sustained-clock leg (0 = skip)")
    ap.add_argument("--workload", default="1080p", choices=sorted(WORKLOADS), help="BASELINE config (default: the metric's)")
    args = ap.parse_args()
    select_workload(args.workload)
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup
    rank = int(os.environ.get("RANK", "0")); local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    from opencv_b200 import HOG
    from opencv_b200.shard import shard_indices, gather_detections

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local_rank)
    binding = bind_near_gpu(local_rank) if world > 1 and os.environ.get("HOGB200_NUMA_BIND", "1") != "0" else "single rank: not bound"
    print("rank %d: %s" % (rank, binding), file=sys.stderr)
    host_group = None
    if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
        os.environ["NCCL_DEBUG"] = "WARN"      # keep NCCL's version banner off stdout: rank 0 prints ONE JSON line
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        host_group = dist.new_group(backend="gloo")        # the host gather of boxes (no collective on the data path)

    CONFIG = make_config()
    hog = HOG.create(device=local_rank)
    people = hog.getDefaultPeopleDetector()
    hog.setSVMDetector(people)
    if NSLOTS > 1:       # the reference ships no vehicle model: fixed-seed stand-in, same geometry
        veh = people + 0.02 * np.random.RandomState(3).standard_normal(3781).astype(np.float32)
        hog.setSVMDetector(veh.astype(np.float32), 1)
    hog.setNumLevels(NLEVELS); hog.setScaleFactor(SCALE0); hog.setHitThreshold(HIT); hog.setGroupThreshold(GROUP)
    stream = torch.cuda.Stream()
    hog.setStream(stream.cuda_stream)

    host = make_frames(BATCH, 1000 + 100 * rank)
    pinned = [torch.from_numpy(f).pin_memory() for f in host]
    dev = [p.cuda(non_blocking=False) for p in pinned]
    pinned_np = [p.numpy() for p in pinned]
    torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # global frame g of a step lives on rank g % world (opencv_b200/shard.py); this rank holds its BATCH frames
    n_global = BATCH * world
    my_ids = shard_indices(n_global, rank, world)

    def finish_step(out, gather):
        if NSLOTS > 1:                             # one list per frame: its detectors' boxes together
            out = [np.concatenate(out[i * NSLOTS:(i + 1) * NSLOTS]) for i in range(len(out) // NSLOTS)]
        t = hog.lastTimings()
        if gather:                                 # boxes of all ranks -> rank 0, frame order, on the host (shard.py)
            n_all = len(out) * world                # (a leg may run a prefix of the step's frames)
            merged = gather_detections(dict(zip(my_ids[:len(out)], out)), n_all, rank, world, group=host_group)
            assert merged is None or len(merged) == n_all
        return sum(len(o) for o in out), t["kernel_launches"], t["candidates"]

    def timed(frames, steps, gather=False, capacity=1 << 20, seconds=None, pipelined=True):
        """K steps (or, with `seconds`, as many steps as fit) between two events on the launching stream.  pipelined: the
        asynchronous entry points with two batches in flight (submit k + 1, then collect k) -- the way a frame loop drives
        the library; else the blocking call."""
        barrier()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        nrects = 0; launches = 0; cands = 0; done = 0
        t0 = time.perf_counter()
        more = lambda: done < steps or (seconds is not None and time.perf_counter() - t0 < seconds)
        with torch.cuda.stream(stream):
            e0.record(stream)
            if pipelined:
                pending = hog.submitBatch(frames, nslots=NSLOTS)
                while True:
                    done += 1
                    nxt = hog.submitBatch(frames, nslots=NSLOTS) if more() else None
                    tw = time.perf_counter()
                    a, b, c = finish_step(hog.collectBatch(pending, capacity=capacity), gather)
                    if os.environ.get("HOGB200_BENCH_VERBOSE"):
                        print("  collect %.1f ms, %d candidates" % (1e3 * (time.perf_counter() - tw), c), file=sys.stderr)
                    nrects += a; launches += b; cands += c
                    if nxt is None:
                        break
                    pending = nxt
            else:
                while more():
                    a, b, c = finish_step(hog.detectMultiScaleBatch(frames, nslots=NSLOTS, capacity=capacity), gather)
                    nrects += a; launches += b; cands += c; done += 1
            e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if world > 1:
            tms = torch.tensor([ms], device="cuda"); dist.all_reduce(tms, op=dist.ReduceOp.MAX); ms = float(tms.item())
        barrier()
        return ms, nrects, launches, cands, done

    # ---- device-resident inputs (value)
    timed(dev, args.warmup)
    sampler = ClockSampler(local_rank).start()
    ms, nrects, launches, _, _ = timed(dev, args.steps)
    clocks = sampler.stop()
    fps = world * BATCH * args.steps / (ms * 1e-3)

    # ---- end to end: pinned host frames through the same public call
    timed(pinned_np, args.warmup, gather=True)
    ms_e2e, nrects_e2e, _, cands_e2e, _ = timed(pinned_np, args.steps, gather=True)
    fps_e2e = world * BATCH * args.steps / (ms_e2e * 1e-3)
    h2d = BATCH * W * H * CN
    nchunks = (BATCH + 63) // 64

    def d2h_bytes(cands_per_step, rects_per_step):
        """device -> host bytes of a step: one 16-byte count per chunk, then either the candidates (16 B each, host tail)
        or -- from 1 024 candidates per chunk, where the tail runs on the device -- the per-segment counts and the rects"""
        if cands_per_step / nchunks >= 1024:
            return int(16 * nchunks + 4 * (BATCH * NSLOTS + nchunks) + 16 * rects_per_step)
        return int(16 * nchunks + 16 * cands_per_step)
    d2h = d2h_bytes(cands_e2e / max(args.steps, 1), nrects_e2e / max(args.steps, 1))
    # the blocking entry point (one batch at a time, every call ends with a synchronisation), for comparison
    timed(dev, 3, pipelined=False)
    ms_b, _, _, _, _ = timed(dev, args.steps, pipelined=False)
    timed(pinned_np, 3, gather=True, pipelined=False)
    ms_be, _, _, _, _ = timed(pinned_np, args.steps, gather=True, pipelined=False)

    line = {"metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": CONFIG, "clocks": clocks,
            "e2e": {"value": fps_e2e, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": launches, "detections_per_step": nrects / max(args.steps, 1),
            "api": "hogb200_submit_batch / hogb200_collect_batch, two batches in flight (value and e2e alike)",
            "blocking_call": {"value": world * BATCH * args.steps / (ms_b * 1e-3), "e2e_value": world * BATCH * args.steps / (ms_be * 1e-3),
                              "unit": UNIT, "api": "hogb200_detect_multiscale_batch"}}

    # ---- the same end-to-end measurement with detections: a hit threshold at which every frame yields thousands of
    # candidates (chosen from frame 0's own score distribution), so compaction under load, the candidate D2H, the
    # ordering, rect scaling and groupRectangles (hog.cpp:471-487) are inside a timed number
    if not args.quick:
        hog.setHitThreshold(-100.0); hog.setGroupThreshold(0)
        _, sc0 = hog.detectMultiScale(dev[0], confidences=True)
        target = min(4000, max(1, len(sc0) // 8))
        thr = float(np.sort(sc0)[-target]) if len(sc0) else 0.0
        hog.setHitThreshold(thr); hog.setGroupThreshold(max(GROUP, 2))
        timed(pinned_np, 3, gather=True)
        ms_h, nrects_h, _, cands_h, _ = timed(pinned_np, args.steps, gather=True)
        timed(dev, 3)
        ms_hr, _, _, _, _ = timed(dev, args.steps)
        line["e2e_with_hits"] = {"value": world * BATCH * args.steps / (ms_h * 1e-3), "unit": UNIT,
                                 "resident_value": world * BATCH * args.steps / (ms_hr * 1e-3),
                                 "hit_threshold": thr, "group_threshold": max(GROUP, 2),
                                 "candidates_per_step": cands_h / args.steps, "candidates_per_frame": cands_h / args.steps / BATCH,
                                 "detections_per_step": nrects_h / args.steps,
                                 "h2d_bytes_per_step": h2d,
                                 "d2h_bytes_per_step": d2h_bytes(cands_h / args.steps, nrects_h / args.steps),
                                 "ms_per_step": ms_h / args.steps}
        hog.setHitThreshold(HIT); hog.setGroupThreshold(GROUP)

    # ---- dense-pyramid stress, "max candidate windows" for real: the stock bias gives ~0
    # hits on synthetic frames, so the run is repeated with a bias that makes every window pass
    if args.workload == "dense" and not args.quick:
        allpass = people.copy(); allpass[-1] = 100.0
        hog.setSVMDetector(allpass)
        nb = min(8, BATCH)
        cap = nb * CONFIG["windows"] + 1024
        # (blocking entry point: it sizes the candidate list to the worst case; the asynchronous one would overflow its
        # 262 144-candidate region and re-run every chunk)
        timed(dev[:nb], 2, capacity=cap, pipelined=False)
        ms_p, nrects_p, _, cands_p, _ = timed(dev[:nb], 5, capacity=cap, pipelined=False)
        timed(pinned_np[:nb], 1, gather=True, capacity=cap, pipelined=False)
        ms_pe, _, _, _, _ = timed(pinned_np[:nb], 5, gather=True, capacity=cap, pipelined=False)
        line["pass_all_bias"] = {"value": world * nb * 5 / (ms_p * 1e-3), "e2e_value": world * nb * 5 / (ms_pe * 1e-3), "unit": UNIT,
                                 "frames_per_step_per_gpu": nb, "candidates_per_frame": cands_p / 5 / nb,
                                 "rects_per_frame": nrects_p / 5 / nb,
                                 "d2h_bytes_per_step": int(16 * cands_p / 5), "ms_per_step": ms_p / 5,
                                 "note": "default weights, bias +100: every window of every level is a candidate"}
        hog.setSVMDetector(people)

    # ---- seconds-long run at settled clocks
    if args.sustain > 0 and not args.quick:
        s2 = ClockSampler(local_rank, 0.05).start()
        ms_s, _, _, _, n_s = timed(dev, 1, seconds=args.sustain)
        line["sustained"] = {"value": world * BATCH * n_s / (ms_s * 1e-3), "unit": UNIT, "seconds": ms_s * 1e-3, "steps": n_s,
                             "clocks": s2.stop()}

    # ---- per-kernel device times (CUDA events on the launching stream, inside the library) for the roofline.
    # The timed region above issues one launch of each kernel per GROUP of frames (4 by default), so the roofline
    # launch is a group launch; the single-frame launch (latency path) is reported next to it.
    hog.enableTiming(True)
    group = hog.batchGroup(W, H, CN)      # frames per launch: the library's choice for this shape (HOGB200_GROUP forces)
    ks = {"resize_ms": [], "hist_ms": [], "svm_ms": [], "total_ms": []}
    ks1 = {"resize_ms": [], "hist_ms": [], "svm_ms": [], "total_ms": []}
    for i in range(3 + 12):
        hog.detectMultiScaleBatch([dev[(i * group + j) % BATCH] for j in range(group)], nslots=NSLOTS, capacity=1 << 20)
        t = hog.lastTimings()
        hog.detectMultiScale(dev[i % BATCH])
        t1 = hog.lastTimings()
        if i >= 3:
            for k in ks:
                ks[k].append(t[k]); ks1[k].append(t1[k])
    hog.enableTiming(False)
    tinfo = hog.lastTimings()
    assert int(tinfo["pyramid_pixels"]) == CONFIG["pyramid_pixels"] and int(tinfo["windows"]) == CONFIG["windows"]
    hist_ms = statistics.mean(ks["hist_ms"])
    alg_bytes = (CN + 2.25) * tinfo["pyramid_pixels"] * group
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    # the committed `ncu --set full` capture of this kernel: quoted only if it was taken from the hist.cu that is running
    traffic, util, issue, cap_note = None, {}, None, "no capture"
    try:
        cap = json.load(open(os.path.join(ROOT, "profiles", "hist_kernel_traffic.json")))
        now = git_blob_hash(os.path.join(ROOT, "opencv_b200", "csrc", "hist.cu"))
        if args.workload != "1080p":
            cap_note = "the committed ncu capture is of the 1080p launch"
        elif cap.get("source_hash", {}).get("opencv_b200/csrc/hist.cu") != now:
            cap_note = "stale: profiles/hist_kernel_traffic.json was captured from another hist.cu (%s)" % str(cap.get("source_hash"))
        else:
            traffic, util, cap_note = cap.get("dram_bytes_per_launch"), cap.get("ncu", {}), "ncu capture of this hist.cu (%s)" % now[:12]
            sm_hz = 1e6 * float(clocks["sm_mhz"] or peaks.get("sm_max_mhz", 1965.0))
            nsm = torch.cuda.get_device_properties(local_rank).multi_processor_count
            slots = nsm * 4 * sm_hz * hist_ms * 1e-3          # warp-instruction issue slots of the launch: SMs x 4 schedulers x clocks
            issue = {"warp_inst_per_launch": util.get("warp_instructions"), "peak_warp_inst_per_launch": slots,
                     "frac": util.get("warp_instructions", 0) / slots if slots else None,
                     "thread_inst_per_pixel": 32 * util.get("warp_instructions", 0) / (tinfo["pyramid_pixels"] * group),
                     "fp32_pipe_pct_ncu": util.get("fma_pipe_pct"), "issue_active_pct_ncu": util.get("issue_active_pct"),
                     "how": "smsp__inst_executed.sum of the capture / (SMs x 4 x median SM clock of this run x live launch time)"}
    except Exception as e:
        cap_note = "no capture (%s)" % e
    achieved = alg_bytes / (hist_ms * 1e-3) / 1e9
    line["roofline"] = {
        "kernel": "hist_kernel<%d> (fused gradient + histogram + L2-Hys, all %d levels of %d frames in one launch)" % (CN, CONFIG["levels"], group),
        "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
        "traffic": traffic, "issue": issue, "capture": cap_note,
        "limiter": "instruction issue + shared-memory bandwidth (DESIGN.md 4.1): HBM would bound it only below ~35 instructions/pixel",
        "ncu": util, "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)",
        "algorithmic_bytes_per_launch": alg_bytes, "launch_ms": hist_ms, "frames_per_launch": group,
        "stage_ms_per_launch": {k: statistics.mean(v) for k, v in ks.items()},
        "stage_ms_single_frame_launch": {k: statistics.mean(v) for k, v in ks1.items()}}

    if rank == 0 and world == 1 and not args.no_cpu:
        r = cpu_reference_fps([host[0]])                      # warm-up
        r = cpu_reference_fps([host[i % BATCH] for i in range(8 if W * H < 4000000 else 4)])
        total = sum(r["times"])
        line["cpu_baseline"] = {"value": len(r["times"]) / total, "unit": UNIT, "cores": r["cores"], "host_cores": host_cores(),
                                "kind": r["kind"],
                                "sample": "%d frames %dx%d of the same workload, stock cv::HOGDescriptor::detectMultiScale%s, "
                                          "best frame %.1f ms" % (len(r["times"]), W, H, " (once per detector)" if NSLOTS > 1 else "",
                                                                  1e3 * min(r["times"]))}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
