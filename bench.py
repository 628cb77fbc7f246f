#!/usr/bin/env python3
"""Benchmark of the ORB feature frontend hot path (BASELINE.json metric: frames/s ORB extract+match @752x480 / 1k feats).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl orbx|reference] [--batch B]

One "step" = one pass of the hot path over one batch of B synthetic 752x480 frames (nFeatures 1000, 8 levels, 1.2, FAST
20/7; SURVEY.md §8-d generator).  Our arm: frames resident in HBM, all kernels of the C-ABI library (liborbx.so) on torch's
current stream, timed with CUDA events on that stream, max over ranks.  torch is plumbing only (HBM buffers, streams,
torch.distributed barrier / max).  Frames shard by rank with no data-path collective ("weak" scaling: B per GPU).

JSON line keys beyond the base contract:
  e2e           same metric through the host-buffer C-ABI call (orbx_extract_batch_host): pinned host frames in, keypoints +
                descriptors out, H2D / D2H inside the timed region
  roofline      dominant kernel of the step (per-stage CUDA events): algorithmic bytes per launch / its duration vs the
                measured HBM peak of MEASURED_PEAKS.json (DESIGN.md §Kernels states the per-kernel byte model)
  cpu_baseline  the CPU oracle (a port of the reference's ORBextractor path; the reference itself cannot be compiled in
                this image) timed on the box's host cores on a bounded sample
  --impl reference  runs only that CPU arm, one bounded sample per step, and prints the same line with impl=reference
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H, NFEAT = 752, 480, 1000
LEVEL_PIXELS = 1117367            # sum over the 8 levels
ALGO_BYTES_PER_FRAME = 1177367    # input + levels 1..7 + N*(28+32)   
METRIC = "frames/s ORB extract+match @752x480/1k feats"
WORKLOAD = "euroc_mono_752x480_nfeat1000_extract"


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """samples nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)"""

    def __init__(self, index):
        self.rows = []
        self.proc = None
        self.index = index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            p = [x.strip() for x in r.split(",")]
            if len(p) < 6:
                continue
            try:
                sm.append(float(p[0])); mx.append(float(p[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_reference_sample(synth, nframes, threads, seed=12345):
    """the reference's CPU path (oracle port) on `nframes` frames of the workload with `threads` host threads -> frames/s"""
    import oracle
    oracle.build()
    frames = synth.frames(W, H, seed, range(nframes))
    sec, kps = oracle.time_extract(frames, threads, nfeatures=NFEAT, blur_taps=oracle.TAPS_OCV346, sincos_mode=oracle.SINCOS_LIBM_F32)
    return nframes / sec, sec, kps


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    synth = importlib.import_module("3_orb_slam3_selfnote_b200.synth")
    cores = os.cpu_count() or 1
    per_step = max(cores, min(4 * cores, 256))
    for _ in range(args.warmup):
        cpu_reference_sample(synth, cores, cores)
    total_sec, total_frames = 0.0, 0
    for s in range(args.steps):
        _, sec, _ = cpu_reference_sample(synth, per_step, cores, seed=1000 + s)
        total_sec += sec
        total_frames += per_step
    fps = total_frames / total_sec
    line = {"metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * total_sec / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic", "impl": "reference",
            "config": {"workload": WORKLOAD, "frames_per_step": per_step, "width": W, "height": H, "nfeatures": NFEAT,
                       "note": "CPU oracle port of the reference ORBextractor path (reference needs OpenCV 3.4 C++: unbuildable here)"},
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port",
                             "sample": "%d frames per step x %d steps, one frame per thread, %d threads" % (per_step, args.steps, cores)},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_orbx(args):
    import torch
    import torch.distributed as dist
    orbx = importlib.import_module("3_orb_slam3_selfnote_b200")
    synth = importlib.import_module("3_orb_slam3_selfnote_b200.synth")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B = args.batch
    # synthetic frames: `distinct` different frames per rank tiled to B (generation is host-side numpy)
    distinct = min(B, 32)
    base = synth.frames(W, H, 1000 + rank, range(distinct))
    host = np.concatenate([base] * ((B + distinct - 1) // distinct))[:B]
    pinned = torch.from_numpy(host).pin_memory()
    dev = pinned.cuda(non_blocking=False)
    gx = orbx.ORBextractor(NFEAT, max_batch=B, device=local)
    stream = torch.cuda.Stream()   # explicit non-default stream: the C-ABI treats a NULL stream as "use the handle's own"
    torch.cuda.set_stream(stream)

    def step():
        gx.extract_batch_device(dev.data_ptr(), B, W, H, W, W * H, (0, 0), stream.cuda_stream)

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.barrier()
        ms = float(t.item())
    clocks = sampler.stop() if rank == 0 else None
    counts, mono, kps, desc = gx.results_host(B)   # also checks the device status word
    kp_per_frame = float(counts.mean())
    value = world * B * args.steps / (ms / 1e3)

    # per-stage device times (profiling pass, outside the timed region)
    gx.set_profiling(True)
    acc = None
    for _ in range(3):
        step()
        torch.cuda.synchronize()
        st = gx.stage_times()
        acc = [(n, t) for n, t in st] if acc is None else [(n, a + t) for (n, a), (_, t) in zip(acc, st)]
    gx.set_profiling(False)
    stages = [(n, t / 3) for n, t in acc]
    # algorithmic bytes per launch of each stage for one frame
    algo = {"k_pyr_resize(x levels-1)": LEVEL_PIXELS + (LEVEL_PIXELS - W * H) - 210 * 134,  # read levels 0..6, write levels 1..7
            "k_fast_nms": LEVEL_PIXELS, "k_blur7": 2 * LEVEL_PIXELS, "k_quadtree": 0, "k_assemble": 0,
            "k_orient_desc": int(kp_per_frame * 60), "memset": 0}
    dom_name, dom_ms = max(stages[1:], key=lambda x: x[1])
    peak, peak_kind = _peaks()
    dom_bytes = algo.get(dom_name, 0) * B
    achieved = dom_bytes / (dom_ms / 1e3) / 1e9 if dom_ms > 0 else 0.0
    step_ms = sum(t for _, t in stages)
    roof = {"bound": "hbm", "kernel": dom_name, "achieved": achieved, "peak": peak, "peak_kind": peak_kind, "unit": "GB/s",
            "frac": achieved / peak, "traffic": None, "kernel_ms_per_launch": dom_ms, "kernel_share_of_step": dom_ms / step_ms,
            "algorithmic_bytes_per_launch": dom_bytes,
            "whole_path": {"achieved": value / world * ALGO_BYTES_PER_FRAME / 1e9, "frac": value / world * ALGO_BYTES_PER_FRAME / 1e9 / peak,
                           "bytes_per_frame": ALGO_BYTES_PER_FRAME},
            "stage_ms": {n: round(t, 4) for n, t in stages}}

    # end to end through the host-buffer C-ABI call: pinned frames in, keypoints + descriptors out
    cap = gx.max_keypoints(W, H)
    okps = torch.empty((B, cap, 28), dtype=torch.uint8).pin_memory()
    odesc = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
    ocnt = torch.empty(B, dtype=torch.int32).pin_memory()
    omono = torch.empty(B, dtype=torch.int32).pin_memory()
    ge = orbx.ORBextractor(NFEAT, max_batch=max(1, B // 4), device=local)   # 4 chunks per call -> copies overlap kernels
    import ctypes as C

    def e2e_step():
        rc = orbx.lib().orbx_extract_batch_host(ge._h, C.c_void_p(pinned.data_ptr()), B, W, H, W, C.c_size_t(W * H), 0, 0,
                                                C.c_void_p(okps.data_ptr()), C.c_void_p(odesc.data_ptr()), cap,
                                                C.c_void_p(ocnt.data_ptr()), C.c_void_p(omono.data_ptr()))
        if rc != 0:
            raise RuntimeError("orbx_extract_batch_host failed: %d %s" % (rc, orbx.lib().orbx_last_error()))

    for _ in range(3):
        e2e_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    assert np.array_equal(ocnt.numpy(), counts), "e2e and device-resident paths disagree"
    e2e = {"value": world * B * args.steps / e2e_s, "unit": "frames/s", "h2d_bytes_per_step": B * W * H,
           "d2h_bytes_per_step": int(B * (cap * 60 + 8)), "timer": "host wall clock around the synchronous C-ABI call"}

    if rank == 0:
        cores = os.cpu_count() or 1
        ncpu = max(cores, min(2 * cores, 128))
        fps1, sec1, _ = cpu_reference_sample(synth, max(8, min(ncpu, 16)), 1)
        fpsN, secN, _ = cpu_reference_sample(synth, ncpu, cores)
        cpu = {"value": fpsN, "unit": "frames/s", "cores": cores, "kind": "port",
               "sample": "%d frames of the workload, one frame per thread on %d threads (%.1f s); single thread: %.1f frames/s"
                         % (ncpu, cores, secN, fps1), "single_thread_value": fps1}
        line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
                "data": "synthetic",
                "config": {"workload": WORKLOAD, "frames_per_step_per_gpu": B, "width": W, "height": H, "nfeatures": NFEAT,
                           "levels": 8, "scale": 1.2, "fast": [20, 7], "keypoints_per_frame": kp_per_frame,
                           "l2": "per-step working set (%.0f MB of frames + pyramids + blurred levels per GPU) exceeds the 126 MB L2"
                                 % (B * (W * H + 2 * LEVEL_PIXELS) / 1e6),
                           "sharding": "frames partitioned by rank, no collective on the data path"},
                "clocks": clocks, "e2e": e2e, "gpu_launches": gx.launches_per_batch() * args.steps * world,
                "roofline": roof, "cpu_baseline": cpu}
        print(json.dumps(line), flush=True)
    gx.close(); ge.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="orbx", choices=["orbx", "reference"])
    ap.add_argument("--batch", type=int, default=256)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_orbx(args)


if __name__ == "__main__":
    main()
