#!/usr/bin/env python3
"""Benchmark of the ORB feature frontend hot path (BASELINE.json metric: frames/s ORB extract+match @752x480).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl orbx|reference] [--batch B]

Workload = BASELINE.json configs[1]: EuRoC-shaped rectified stereo, synthetic 752x480 left/right pairs (SURVEY.md §8-d
generator), nFeatures 1200 per image, 8 levels, 1.2, FAST 20/7, extraction of both images + Frame::ComputeStereoMatches
(row-constrained Hamming search, 11x11 SAD refinement, median cut).  One "step" = one batch of B stereo frames; a "frame"
is one stereo frame (= 2 images), as in the reference's Frame object.

Our arm: frames resident in HBM, every kernel of the C-ABI library (liborbx.so) enqueued on one CUDA stream, timed with
CUDA events on that stream, max over ranks.  torch is plumbing only (HBM buffers, the stream, torch.distributed barrier /
max).  Stereo frames shard by rank with no data-path collective ("weak" scaling: B frames per GPU).

JSON keys beyond the base contract:
  e2e           the same metric through the host-buffer C-ABI call orbx_stereo_frames_host: pinned host images in; keypoints,
                descriptors, uRight and depth out; H2D / D2H inside the timed region
  roofline      dominant kernel of the step (per-stage CUDA events of the extractor): algorithmic bytes per launch / its
                duration vs the measured HBM peak (MEASURED_PEAKS.json); DESIGN.md states the per-kernel byte model
  cpu_baseline  the CPU oracle (port of the reference's ORBextractor + ComputeStereoMatches; the reference itself needs
                OpenCV 3.4 C++ and cannot be compiled in this image) on the box's host cores, bounded sample
  extra         mono extraction-only throughput (752x480, nFeatures 1000: the shape the north-star target is quoted on) and
                informational device-resident runs of the other BASELINE.json configs (HD shapes, TUM-VI shape with
                SearchByProjection, CLAHE / remap / cvtColor in front of the extractor); --no-extras skips them
  --impl reference  runs only the CPU arm, one bounded sample per step, same line with impl=reference
"""
import argparse
import ctypes as C
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

W, H, NFEAT = 752, 480, 1200
LEVEL_PIXELS = 1117367            # sum over the 8 levels
MBF, FX = np.float32(47.90639384423901), np.float32(435.2046959714599)   # Examples/Stereo/EuRoC.yaml:9,28
MB = np.float32(MBF / FX)
METRIC = "frames/s ORB extract+match @752x480 (EuRoC-shaped stereo pairs, 1200 feats/image)"
WORKLOAD = "euroc_stereo_752x480_nfeat1200_extractLR+ComputeStereoMatches"


def algo_bytes_per_stereo_frame(kp_per_image):
    # per image: read input + write levels 1..7 + keypoints/descriptors; per frame: + uRight/depth
    per_image = W * H + (LEVEL_PIXELS - W * H) + kp_per_image * 60
    return 2 * per_image + kp_per_image * 8


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


NCU_SUMMARY = os.path.join(ROOT, "profiles", "r02m_ncu_full_summary.csv")   # tools/ncu_summary.py of one --set full capture (128 images per launch)
NCU_NAME = {"k_fast_nms": "k_fast_nms3", "k_blur7": "k_blur7v", "k_orient_desc": "k_orient_desc2", "k_quadtree": "k_quadtree"}


def _kernel_is(ncu_name, name):
    """ncu prints templated kernels as `void name<args>(...)`, plain ones as `name(...)`"""
    base = ncu_name.split("(")[0].replace("void ", "").split("<")[0].strip()
    return base == name


def _ncu_traffic(stage, images_per_launch):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the stage's kernel from the committed ncu capture, or None.
    The capture ran 128 images per launch; the bench launch is scaled linearly when its batch differs."""
    import csv
    try:
        rows = list(csv.reader(open(NCU_SUMMARY)))
        hdr, units = rows[0], rows[1]
        ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        name = NCU_NAME.get(stage)
        for r in rows[2:]:
            if name and _kernel_is(r[0], name):
                return (float(r[ir]) * mult[units[ir]] + float(r[iw]) * mult[units[iw]]) * images_per_launch / 128.0
    except Exception:
        pass
    return None


class ClockSampler:
    """samples nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)"""

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                                          "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
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


def cv2_primitive_times(synth, nframes=6):
    """the real OpenCV's (python cv2, one thread) time for the OpenCV primitives of one 752x480 stereo frame
    -- resize chain, FAST on every level (whole level, not per cell: a lower bound), 7x7 GaussianBlur, BFMatcher 2-NN of
    1200 x 1200 descriptors -- as a lower bound for those stages of the reference (the scalar oracle port is slower than
    OpenCV's SIMD).  Reported only; quadtree, orientation, rBRIEF and the stereo search are not in it."""
    try:
        import cv2
    except Exception:
        return None
    cv2.setNumThreads(1)
    imgs = [synth.frame(W, H, 4242, i) for i in range(nframes)]
    sizes = [(W, H)]
    for _ in range(7):
        sizes.append((int(round(sizes[-1][0] / 1.2)), int(round(sizes[-1][1] / 1.2))))   # close enough for a timing
    fast = cv2.FastFeatureDetector_create(20, True)
    t = {"resize_chain": 0.0, "fast_levels": 0.0, "blur_levels": 0.0}
    for img in imgs:
        t0 = time.perf_counter()
        pyr = [img]
        for (w_, h_) in sizes[1:]:
            pyr.append(cv2.resize(pyr[-1], (w_, h_), interpolation=cv2.INTER_LINEAR))
        t1 = time.perf_counter()
        for lv in pyr:
            fast.detect(lv, None)
        t2 = time.perf_counter()
        for lv in pyr:
            cv2.GaussianBlur(lv, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
        t3 = time.perf_counter()
        t["resize_chain"] += t1 - t0; t["fast_levels"] += t2 - t1; t["blur_levels"] += t3 - t2
    rng = np.random.default_rng(0)
    q = rng.integers(0, 256, (1200, 32), dtype=np.uint8); tr = rng.integers(0, 256, (1200, 32), dtype=np.uint8)
    bf = cv2.BFMatcher(cv2.NORM_HAMMING)
    t0 = time.perf_counter()
    for _ in range(3):
        bf.knnMatch(q, tr, k=2)
    knn = (time.perf_counter() - t0) / 3
    per_image = {k: v / nframes * 1e3 for k, v in t.items()}
    return {"threads": 1, "cv2": cv2.__version__, "ms_per_image": per_image, "bf_knn2_1200x1200_ms": knn * 1e3,
            "ms_per_stereo_frame_lower_bound": 2 * sum(per_image.values())}


def static_config():
    """the workload description, IDENTICAL in both arms (the driver compares the two `config` objects)"""
    return {"workload": WORKLOAD, "images_per_frame": 2, "width": W, "height": H, "nfeatures": NFEAT, "levels": 8, "scale": 1.2, "fast": [20, 7],
            "frame": "one rectified stereo pair: ORBextractor on both images + Frame::ComputeStereoMatches",
            "l2": "inputs larger than L2: every launch reads a batch of images + pyramids + blurred levels far above the 126 MB L2",
            "sharding": "stereo frames partitioned by rank, no collective on the data path"}


def cpu_stereo_sample(synth, nframes, threads, seed=12345):
    """the reference's CPU path on `nframes` stereo frames with `threads` host threads -> (frames/s, seconds, matched, kind).
    kind "reference": the reference's OWN src/ORBextractor.cc + Frame::ComputeStereoMatches compiled unmodified (oracle/_ref,
    prebuilt where /root/reference exists); kind "port": the oracle restatement when that library is absent."""
    import oracle
    from oracle import refmatch
    left = synth.frames(W, H, seed, range(nframes), False)
    right = synth.frames(W, H, seed, range(nframes), True)
    if refmatch.available():
        sec, matched = refmatch.time_stereo(left, right, threads, NFEAT, float(MB), float(MBF))
        return nframes / sec, sec, matched, "reference"
    oracle.build()
    sec, matched = oracle.time_stereo(left, right, threads, NFEAT, float(MB), float(MBF))
    return nframes / sec, sec, matched, "port"


CPU_NOTE = ("oracle/_ref: the reference's unmodified ORBextractor.cc + ComputeStereoMatches built -O3 -march=x86-64-v3 against stand-in cv:: "
            "primitives (scalar restatements of OpenCV's resize / FAST / GaussianBlur, slower than OpenCV's SIMD: see cv2_primitives for the floor)")


def run_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    synth = importlib.import_module("3_orb_slam3_selfnote_b200.synth")
    cores = os.cpu_count() or 1
    per_step = max(cores // 2, min(2 * cores, 128))
    kind = "port"
    for _ in range(min(args.warmup, 2)):
        cpu_stereo_sample(synth, max(2, cores // 2), cores)
    total_sec, total_frames = 0.0, 0
    for s_ in range(args.steps):
        _, sec, _, kind = cpu_stereo_sample(synth, per_step, cores, seed=1000 + s_)
        total_sec += sec
        total_frames += per_step
    fps = total_frames / total_sec
    sample = "%d stereo frames per step x %d steps; frame-parallel, 2 threads per frame (Frame.cc:120-121), %d threads" % (per_step, args.steps, cores)
    line = {"metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * total_sec / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic", "impl": "reference", "config": static_config(),
            "run": {"frames_per_step": per_step, "note": CPU_NOTE if kind == "reference" else "oracle port (oracle/_ref not built)"},
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    _emit(line)


def extra_configs(orbx, synth, torch, stream, local, steps):
    """Informational device-resident runs of the other BASELINE.json configs on one GPU (the parity of each is covered by tests/):
    config 2 (TUM-VI shaped 512x512 mono, 1500 features, extraction + SearchByProjection against 5000 projected map points per
    frame), config 3 (1280x720 and 1920x1080, 2000 features, 256 frames per batch) and the tile kernel of config 4's all-pairs
    keyframe 2-NN sweep (descriptor pairs/s against the popc roofline of SURVEY.md section 8-d)."""
    sh = stream.cuda_stream
    out = {}

    def timed(fn, reps):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / 1e3 / reps

    def tiled(w, h, n, distinct, seed):
        fr = synth.frames(w, h, seed, range(distinct), False)
        return torch.from_numpy(np.concatenate([fr] * ((n + distinct - 1) // distinct))[:n]).cuda()

    # HD shapes, 2000 features, 256 frames per batch
    for (w, h, name) in ((1280, 720, "hd720"), (1920, 1080, "hd1080")):
        n = 256
        dev = tiled(w, h, n, 4, 3000)
        ex = orbx.ORBextractor(2000, max_batch=n, device=local)
        sec = timed(lambda: ex.extract_batch_device(dev.data_ptr(), n, w, h, w, w * h, (0, 0), sh), steps)
        counts = ex.results_host(n)[0]
        out["%s_extract_nfeat2000_batch256_frames_per_s_per_gpu" % name] = n / sec
        out["%s_keypoints_per_frame" % name] = float(counts.mean())
        ex.close()
        del dev
    # TUM-VI shape: extraction + windowed matching against 5000 projected map points per frame (host map-point arrays in, matches out)
    w = h = 512
    n, nmp = 64, 5000
    dev = tiled(w, h, n, 8, 4000)
    ex = orbx.ORBextractor(1500, max_batch=n, device=local)
    mt = orbx.ORBmatcher(0.8, device=local)
    ex.extract_batch_device(dev.data_ptr(), n, w, h, w, w * h, (0, 0), sh)
    torch.cuda.synchronize()
    counts, _, kps, desc = ex.results_host(n)
    rng = np.random.default_rng(7)
    px = np.empty((n, nmp), np.float32); py = np.empty((n, nmp), np.float32); lv = np.empty((n, nmp), np.int32)
    md = rng.integers(0, 256, (n, nmp, 32), dtype=np.uint8)
    for f in range(n):   # map points = this frame's own keypoints shifted by a pixel + uniformly random padding
        c = int(counts[f])
        px[f, :c] = kps[f, :c]["x"] + 1.0; py[f, :c] = kps[f, :c]["y"] - 1.0; lv[f, :c] = kps[f, :c]["octave"]; md[f, :c] = desc[f, :c]
        px[f, c:] = rng.uniform(0, w, nmp - c); py[f, c:] = rng.uniform(0, h, nmp - c); lv[f, c:] = rng.integers(0, 8, nmp - c)
    inv = np.ones((n, nmp), np.uint8)
    vc = np.where(np.arange(nmp) % 2 == 0, 1.0, 0.9).astype(np.float32)[None].repeat(n, 0)
    nm_holder = {}

    def tum_step():
        ex.extract_batch_device(dev.data_ptr(), n, w, h, w, w * h, (0, 0), sh)
        nm_holder["nm"] = mt.SearchByProjectionBatch(ex, n, inv, px, py, lv, vc, md, 1.0)[0]

    ex._last_shape = (h, w)
    sec = timed(tum_step, max(2, steps // 2))
    out["tumvi512_extract_nfeat1500+SearchByProjection_5000mp_frames_per_s_per_gpu"] = n / sec
    out["tumvi512_matches_per_frame"] = float(np.mean(nm_holder["nm"]))
    # the same shape through the device-resident path: the 5000-point local map stays in HBM, per frame only
    # the pose goes in; isInFrustum + SearchByProjection run on the extractor's stream, kp_mp / nmatches come back once per batch
    pre_mod = importlib.import_module("3_orb_slam3_selfnote_b200.preproc")
    fx, fy, cx, cy = np.float32(190.978), np.float32(190.973), np.float32(254.93), np.float32(256.9)
    kb = (0.003482389402, 0.000715034845, -0.002053236141, 0.000202936736)   # Examples/Monocular/TUM_512.yaml: KannalaBrandt8 k1..k4
    zs = rng.uniform(1.0, 8.0, nmp)
    mx, my = (px[0].astype(np.float64) - cx) / fx, (py[0].astype(np.float64) - cy) / fy   # KannalaBrandt8::unproject by Newton steps
    rr = np.maximum(np.hypot(mx, my), 1e-9)
    th = rr.copy()
    for _ in range(8):
        t2 = th * th
        f_ = th * (1 + t2 * (kb[0] + t2 * (kb[1] + t2 * (kb[2] + t2 * kb[3])))) - rr
        df = 1 + t2 * (3 * kb[0] + t2 * (5 * kb[1] + t2 * (7 * kb[2] + t2 * 9 * kb[3])))
        th = th - f_ / df
    P = (np.stack([np.sin(th) * mx / rr, np.sin(th) * my / rr, np.cos(th)], 1) * zs[:, None]).astype(np.float32)   # frame 0's map points, camera at the origin
    dist = np.linalg.norm(P, axis=1)
    scale = ex.GetScaleFactors()
    maxd = (dist * scale[np.clip(lv[0], 0, 7)] * 0.95).astype(np.float32)
    mind = (maxd / scale[7] * 0.8).astype(np.float32)
    nrm = (P / dist[:, None]).astype(np.float32)
    lm = orbx.LocalMap(mt, P, nrm, maxd, mind, md[0])
    cams = [pre_mod.frustum(np.eye(3), np.zeros(3), fx, fy, cx, cy, (0, w, 0, h), 0.0, 1.2, 8, kb=kb) for _ in range(n)]

    def tum_step_dev():
        ex.extract_batch_device(dev.data_ptr(), n, w, h, w, w * h, (0, 0), sh)
        nm_holder["nm_dev"] = mt.SearchLocalPointsBatch(ex, n, cams, lm, 1.0, sh)[0]

    sec = timed(tum_step_dev, max(2, steps // 2))
    out["tumvi512_extract_nfeat1500+SearchLocalPoints_5000mp_device_map_frames_per_s_per_gpu"] = n / sec
    out["tumvi512_device_map_matches_per_frame"] = float(np.mean(nm_holder["nm_dev"]))
    # ... and with the CLAHE(3.0, 8x8) that mono_tum_vi.cc:109 applies to every frame in front of it
    pre_t = pre_mod.Preprocessor(device=local)
    dev_eq = torch.empty_like(dev)

    def tum_step_clahe():
        pre_t.clahe_device(dev.data_ptr(), n, w, h, w, w * h, dev_eq.data_ptr(), w, w * h, 3.0, (8, 8), sh)
        ex.extract_batch_device(dev_eq.data_ptr(), n, w, h, w, w * h, (0, 0), sh)
        nm_holder["nm_eq"] = mt.SearchLocalPointsBatch(ex, n, cams, lm, 1.0, sh)[0]

    sec = timed(tum_step_clahe, max(2, steps // 2))
    out["tumvi512_clahe+extract_nfeat1500+SearchLocalPoints_5000mp_frames_per_s_per_gpu"] = n / sec
    pre_t.close()
    lm.close()
    ex.close(); mt.close()
    # pre-processing in front of the extractor, device-resident, 256 frames of 752x480 per call
    pre = pre_mod.Preprocessor(device=local)
    nb = 256
    g8 = tiled(W, H, nb, 4, 5000)
    col = torch.randint(0, 256, (nb, H, W, 3), dtype=torch.uint8, device="cuda")
    dst = torch.empty_like(g8)
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
    r2 = ((xx - W / 2) ** 2 + (yy - H / 2) ** 2) / (0.6 * W) ** 2
    kk = 1 - 0.28 * r2 + 0.07 * r2 * r2
    tbl = pre.remap_table(((xx - W / 2) * kk * 1.08 + W / 2).astype(np.float32), ((yy - H / 2) * kk * 1.08 + H / 2).astype(np.float32), (W, H))
    sec = timed(lambda: pre.cvt_gray_device(col.data_ptr(), nb, W, H, W * 3, H * W * 3, 3, False, dst.data_ptr(), W, H * W, sh), steps)
    out["cvtColor_bgr2gray_752x480_frames_per_s_per_gpu"] = nb / sec
    out["cvtColor_bgr2gray_GBps"] = nb * H * W * 4 / sec / 1e9
    sec = timed(lambda: pre.remap_device(tbl, g8.data_ptr(), nb, W, H * W, dst.data_ptr(), W, H * W, sh), steps)
    out["remap_rectify_752x480_frames_per_s_per_gpu"] = nb / sec
    sec = timed(lambda: pre.clahe_device(g8.data_ptr(), nb, W, H, W, H * W, dst.data_ptr(), W, H * W, 3.0, (8, 8), sh), steps)
    out["clahe_3.0_8x8_752x480_frames_per_s_per_gpu"] = nb / sec
    tbl.close(); pre.close()
    del col, g8, dst
    return out


def _bind_to_gpu_numa_node(torch, local):
    """Multi-rank runs: pin this rank (and the e2e worker threads it starts) to the CPUs of the NUMA node its GPU hangs off, so
    the page-locked staging buffers are allocated next to the GPU's PCIe root.  Returns the cpulist used, or None."""
    try:
        p = torch.cuda.get_device_properties(local)
        bus = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        with open("/sys/bus/pci/devices/%s/local_cpulist" % bus) as f:
            cl = f.read().strip()
        cpus = set()
        for part in cl.split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        cpus &= os.sched_getaffinity(0)
        if len(cpus) < 2:
            return None
        os.sched_setaffinity(0, cpus)
        return cl
    except Exception:
        return None


STEREO_KERNELS = 4   # k_stereo_rowtable, k_stereo_hamming2, k_stereo_sad, k_stereo_median
NCU_ISSUE = NCU_SUMMARY   # the same capture: smsp__inst_executed.sum per launch


def _popc_roof_pairs_per_s():
    """issue roof of the 256-bit Hamming pair: 8 POPC per pair on a pipe that retires 16 POPC/clk/SM (tools/microbench/popc_rate.cu
    measures 15.0 of them on this pool's B200s with the loop overhead of the probe, profiles/r02_popc_rate.txt): 16 x 148 SMs x SM clock / 8"""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            mhz = float(json.load(f).get("sm_max_mhz", 1965.0))
    except Exception:
        mhz = 1965.0
    measured = None
    try:
        for ln in open(os.path.join(ROOT, "profiles", "r02_popc_rate.txt")):
            if ln.startswith("pair: 8 x"):
                measured = float(ln.split()[-2])
    except Exception:
        pass
    return 16 * 148 * mhz * 1e6 / 8, measured


def _ncu_issue(stage, images_per_launch, kernel_ms, sm_mhz):
    """issue-slot utilisation of the stage's kernel: warp instructions per launch (smsp__inst_executed.sum of the committed ncu capture,
    scaled to this launch's image count) / (4 schedulers x 148 SMs x cycles of the LIVE kernel duration)"""
    import csv
    try:
        rows = list(csv.reader(open(NCU_ISSUE)))
        hdr = rows[0]
        ii = hdr.index("smsp__inst_executed.sum")
        ib = hdr.index("images_per_launch") if "images_per_launch" in hdr else None
        name = NCU_NAME.get(stage)
        for r in rows[2:]:
            if name and _kernel_is(r[0], name):
                per = float(r[ib]) if ib is not None else 128.0
                inst = float(r[ii]) * images_per_launch / per
                cyc = kernel_ms * 1e-3 * sm_mhz * 1e6
                return {"warp_inst_per_launch": inst, "frac": inst / (cyc * 4 * 148), "source": os.path.relpath(NCU_ISSUE, ROOT)}
    except Exception:
        pass
    return None


def allpairs_block(orbx, torch, dist, mt, rank, world, local, kf_total, desc_per_kf, nchunks):
    """BASELINE.json configs[4], CI variant of SURVEY.md 8-d: all-pairs keyframe-descriptor 2-NN over kf_total keyframes x desc_per_kf
    descriptors sharded over the ranks; orbx_allpairs_knn2 (ncclAllGather chunks on a side stream + k_knn2 folds).  Timed once after a
    small warm-up sweep; device time = max over ranks of the library's own event timing (first fold start .. last fold end)."""
    ap = importlib.import_module("3_orb_slam3_selfnote_b200.allpairs")
    mmod = importlib.import_module("3_orb_slam3_selfnote_b200.matcher")
    kf_rank = kf_total // world
    rows = kf_rank * desc_per_kf
    g = torch.Generator(device="cuda")
    g.manual_seed(777 + rank)
    desc = torch.randint(0, 256, (rows, 32), dtype=torch.uint8, device="cuda", generator=g)
    comm = mmod.nccl_comm_from_torch(local) if world > 1 else None
    try:
        ap.allpairs_knn2(desc[: 40 * desc_per_kf].contiguous(), desc_per_kf, matcher=mt, nchunks=2, comm=comm)     # warm-up (NCCL channels, kernels)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        idx, dst, st = ap.allpairs_knn2(desc, desc_per_kf, matcher=mt, nchunks=nchunks, comm=comm, return_stats=True)
        wall = time.perf_counter() - t0
    finally:
        if comm is not None:
            comm.close()
    vals = torch.tensor([st["total_ms"], st["fold_ms"], st["gather_ms"], wall * 1e3], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(vals, op=dist.ReduceOp.MAX)
    total_ms, fold_ms, gather_ms, wall_ms = [float(x) for x in vals.tolist()]
    # spot check (outside the timed region): 16 of this rank's rows against a brute-force scan of ALL rows on the device
    if world > 1:
        allrows = torch.empty((world * rows, 32), dtype=torch.uint8, device="cuda")
        dist.all_gather_into_tensor(allrows, desc)
    else:
        allrows = desc
    lut = torch.tensor([bin(i).count("1") for i in range(256)], dtype=torch.int32, device="cuda")
    ar = torch.arange(world * rows, device="cuda", dtype=torch.int64)
    rng = np.random.default_rng(5 + rank)
    for r in rng.integers(0, rows, 16).tolist():
        gr = rank * rows + r
        d = torch.zeros(world * rows, dtype=torch.int32, device="cuda")
        for c0 in range(0, 32, 8):
            d += lut[(allrows[:, c0:c0 + 8] ^ allrows[gr, c0:c0 + 8]).long()].sum(1, dtype=torch.int32)
        kf0 = (gr // desc_per_kf) * desc_per_kf
        key = (d.long() << 32) | ar
        key[kf0:kf0 + desc_per_kf] = 1 << 62
        top = torch.topk(key, 2, largest=False).values.cpu().numpy()
        want_idx, want_d = top & 0xFFFFFFFF, top >> 32
        if not (np.array_equal(idx[r], want_idx) and np.array_equal(dst[r], want_d)):
            raise AssertionError("allpairs: row %d of rank %d: got %s / %s, brute force %s / %s" % (r, rank, idx[r], dst[r], want_idx, want_d))
    total_pairs = float(world) * rows * (world * rows - desc_per_kf)
    roof, roof_probe = _popc_roof_pairs_per_s()
    pps_gpu = total_pairs / world / (total_ms / 1e3)
    # Which kernel folded: chunks of >= 2^26 pairs go to k_knn2_tc (tcgen05 kind::i8 on +-1 expanded bits: 256 int8 MACs = 512 integer
    # operations per pair), smaller ones to the POPC kernel (orbx_knn2_chunk_device).  Tensor roof: dense int8 is twice the dense bf16 rate
    # on this part (4.5 vs 2.25 POP/s nominal), so 2 x the cuBLAS bf16 throughput measured on this pool (MEASURED_PEAKS.json) is the yardstick.
    chunk_pairs = float(rows) * (rows / max(int(st["nchunks"]), 1))
    tensor = chunk_pairs >= float(1 << 26)
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            bf16_tf = float(json.load(f).get("bf16_tflops", 0.0)) or None
    except Exception:
        bf16_tf = None
    i8_peak = 2.0 * bf16_tf if bf16_tf else 4500.0
    tops = pps_gpu * 512.0 / 1e12
    fold_name = "tensor pipe (k_knn2_tc fold)" if tensor else "popc issue (k_knn2 fold)"
    return {"workload": "allpairs_knn2_%dkf_x_%ddesc" % (kf_total, desc_per_kf), "keyframes": kf_total, "desc_per_kf": desc_per_kf,
            "comm_nranks_seen": int(st["world"]), "nchunks": int(st["nchunks"]), "pairs": total_pairs, "device_s": total_ms / 1e3, "wall_s": wall_ms / 1e3,
            "pairs_per_s": total_pairs / (total_ms / 1e3), "pairs_per_s_per_gpu": pps_gpu,
            "fold_kernel": "k_knn2_tc (tcgen05.mma kind::i8, TMEM accumulators, bulk-copy operand ring)" if tensor else "k_knn2 (xor + popc)",
            "roofline": {"bound": "tensor", "achieved": tops, "peak": i8_peak, "unit": "TOP/s (int8)", "frac": tops / i8_peak,
                         "ops_per_pair": 512, "peak_kind": ("2 x measured dense bf16 (MEASURED_PEAKS.json)" if bf16_tf else "nominal dense int8"),
                         "frac_of_nominal_4500": tops / 4500.0, "ncu": "profiles/r02_knn2_tc.md (tensor pipe active, SASS excerpt)"} if tensor else None,
            "popc_roof_pairs_per_s_per_gpu": roof, "popc_roof_kind": "16 POPC/clk/SM x 148 SMs x max SM clock / 8 POPC per pair (what the xor+popc kernel is bound by)",
            "popc_probe_pairs_per_s_per_gpu": roof_probe, "x_popc_roof": pps_gpu / roof,
            "fold_ms": fold_ms, "allgather_ms": gather_ms, "allgather_bytes_per_rank": float(st["gather_bytes"]), "overlap_frac": float(st["overlap_frac"]),
            "allgather_gbs_per_rank": (float(st["gather_bytes"]) / (gather_ms / 1e3) / 1e9) if gather_ms > 0 else None,
            "limiter": fold_name if fold_ms >= gather_ms else "NVLink all-gather",
            "verified": "16 rows per rank equal a brute-force scan of all rows (index and distance, ties to the lower global row)"}


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
    affinity = _bind_to_gpu_numa_node(torch, local) if world > 1 else None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B, S = args.batch, args.substeps          # frames per launch, launches per step: a step is S*B stereo frames per GPU
    POOL = 2                                  # distinct device-resident sub-batches the launches cycle through
    distinct = min(B, 32)                     # distinct synthetic pairs per sub-batch, tiled to B (generation is host-side numpy)
    reps = (B + distinct - 1) // distinct
    hostL = [np.concatenate([synth.frames(W, H, 1000 + 17 * rank + 7919 * p, range(distinct), False)] * reps)[:B] for p in range(POOL)]
    hostR = [np.concatenate([synth.frames(W, H, 1000 + 17 * rank + 7919 * p, range(distinct), True)] * reps)[:B] for p in range(POOL)]
    devL = [torch.from_numpy(x).cuda() for x in hostL]
    devR = [torch.from_numpy(x).cuda() for x in hostR]
    exL = orbx.ORBextractor(NFEAT, max_batch=B, device=local)
    exR = orbx.ORBextractor(NFEAT, max_batch=B, device=local)
    mt = orbx.ORBmatcher(device=local)
    stream = torch.cuda.Stream()   # explicit non-default stream: the C-ABI treats a NULL stream as "use the handle's own"
    torch.cuda.set_stream(stream)
    sh = stream.cuda_stream

    # the two extractor objects run concurrently, as the reference's two threads do (src/Frame.cc:120-123): the right image goes
    # on a second stream; the stereo matcher joins them.  Timing events sit on `stream`, which waits for the right branch.
    streamR = torch.cuda.Stream()
    shR = streamR.cuda_stream
    evFork, evJoin = torch.cuda.Event(), torch.cuda.Event()

    def launch(p):
        evFork.record(stream)
        streamR.wait_event(evFork)
        exR.extract_batch_device(devR[p].data_ptr(), B, W, H, W, W * H, (0, 0), shR)
        evJoin.record(streamR)
        exL.extract_batch_device(devL[p].data_ptr(), B, W, H, W, W * H, (0, 0), sh)
        stream.wait_event(evJoin)
        mt.stereo_match_batch(exL, exR, MB, MBF, sh)

    def step():
        for k in range(S):
            launch(k % POOL)

    nwarm = max(args.warmup, 3)
    for _ in range(nwarm):
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
    launch(0)                     # results of sub-batch 0 (what the e2e call below processes), outside the timed region
    torch.cuda.synchronize()
    countsL, _, _, _ = exL.results_host(B)   # also checks the device status word
    countsR, _, _, _ = exR.results_host(B)
    uR, depth, _, _ = mt.stereo_results(B)
    kp_per_image = float((countsL.mean() + countsR.mean()) / 2)
    matched = float((uR >= 0).sum() / B)
    value = world * B * S * args.steps / (ms / 1e3)

    # per-stage device times (profiling pass, outside the timed region)
    exL.set_profiling(True)
    acc, stereo_ms = None, 0.0
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        exL.extract_batch_device(devL[0].data_ptr(), B, W, H, W, W * H, (0, 0), sh)
        exR.extract_batch_device(devR[0].data_ptr(), B, W, H, W, W * H, (0, 0), sh)
        s0.record(stream)
        mt.stereo_match_batch(exL, exR, MB, MBF, sh)
        s1.record(stream)
        torch.cuda.synchronize()
        stereo_ms += s0.elapsed_time(s1) / 3
        st = exL.stage_times()
        acc = list(st) if acc is None else [(n, a + t) for (n, a), (_, t) in zip(acc, st)]
    exL.set_profiling(False)
    stages = [(n, t / 3) for n, t in acc]          # one image's extractor stages (a launch runs them twice)
    algo = {"k_pyr_resize(x levels-1)": (LEVEL_PIXELS - 210 * 134) + (LEVEL_PIXELS - W * H),  # read levels 0..6, write levels 1..7
            "k_fast_nms": LEVEL_PIXELS, "k_blur7": 2 * LEVEL_PIXELS, "k_quadtree": 0, "k_assemble": 0,
            "k_orient_desc": int(kp_per_image * 60), "memset": 0}
    dom_name, dom_ms = max(stages[1:], key=lambda x: x[1])
    peak, peak_kind = _peaks()
    dom_bytes = algo.get(dom_name, 0) * B
    achieved = dom_bytes / (dom_ms / 1e3) / 1e9 if dom_ms > 0 else 0.0
    launch_ms = 2 * sum(t for _, t in stages) + stereo_ms
    abf = algo_bytes_per_stereo_frame(kp_per_image)
    sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
    issue = _ncu_issue(dom_name, B, dom_ms, sm_mhz)
    hbm_frac = achieved / peak
    roof = {"bound": "issue" if (issue and issue["frac"] > hbm_frac) else "hbm", "kernel": dom_name, "achieved": achieved, "peak": peak,
            "peak_kind": peak_kind, "unit": "GB/s", "frac": hbm_frac, "issue": issue,
            "traffic": _ncu_traffic(dom_name, B), "traffic_source": os.path.relpath(NCU_SUMMARY, ROOT), "kernel_ms_per_launch": dom_ms,
            "launches_per_step": 2 * S, "kernel_share_of_step": 2 * dom_ms / launch_ms, "algorithmic_bytes_per_launch": dom_bytes,
            "whole_path": {"achieved": value / world * abf / 1e9, "frac": value / world * abf / 1e9 / peak, "bytes_per_frame": abf},
            "stage_ms_per_image_batch": {n: round(t, 4) for n, t in stages}, "stereo_match_ms": round(stereo_ms, 4),
            "note": "bound = the larger of the HBM fraction (algorithmic bytes / live duration / measured copy peak) and the issue-slot "
                    "fraction (ncu warp instructions / live cycles / 592 schedulers); this path is integer ALU work, not a streaming copy"}

    # end to end through the host-buffer C-ABI call: pinned images in; keypoints, descriptors, uRight, depth out
    cap = exL.max_keypoints(W, H)
    nb = max(1, B // args.e2e_chunks)   # chunks per call -> copies overlap kernels
    # host buffers of the e2e call come from orbx_host_alloc (cudaHostAlloc): torch's pin_memory() buffers upload at half the link rate here
    hb = lambda shape, dt: orbx.HostBuffer(shape, dt)
    hinL, hinR = hb((B, H, W), np.uint8), hb((B, H, W), np.uint8)
    hinL.array[...] = hostL[0]
    hinR.array[...] = hostR[0]
    vp = lambda t: t.ptr
    right_records = not args.e2e_no_right

    class Worker:
        """one host thread of the application: its own extractor / matcher handles (handles are independent, include/orbx.h) and its
        own output buffers; issues the synchronous host-buffer call for the sub-batches it is given"""

        def __init__(self):
            self.eL = orbx.ORBextractor(NFEAT, max_batch=nb, device=local)
            self.eR = orbx.ORBextractor(NFEAT, max_batch=nb, device=local)
            self.em = orbx.ORBmatcher(device=local)
            self.okL, self.odL = hb((B, cap, 28), np.uint8), hb((B, cap, 32), np.uint8)
            self.okR = hb((B, cap, 28), np.uint8) if right_records else None
            self.odR = hb((B, cap, 32), np.uint8) if right_records else None
            self.ocL, self.ocR = hb((B,), np.int32), hb((B,), np.int32)
            self.ouR, self.odp = hb((B, cap), np.float32), hb((B, cap), np.float32)

        def call(self):
            rc = orbx.lib().orbx_stereo_frames_host(self.em._m, self.eL._h, self.eR._h, vp(hinL), vp(hinR), B, W, H, W, C.c_size_t(W * H),
                                                    C.c_float(MB), C.c_float(MBF), vp(self.okL), vp(self.odL), vp(self.ocL),
                                                    vp(self.okR) if right_records else None, vp(self.odR) if right_records else None,
                                                    vp(self.ocR), vp(self.ouR), vp(self.odp), cap)
            if rc != 0:
                raise RuntimeError("orbx_stereo_frames_host failed: %d %s" % (rc, orbx.lib().orbx_last_error()))

        def buffers(self):
            return [o for o in (self.eL, self.eR, self.em, self.okL, self.okR, self.odL, self.odR, self.ocL, self.ocR, self.ouR, self.odp) if o is not None]

    nworkers = args.e2e_workers if args.e2e_workers > 0 else (2 if world < 4 else 1)   # 8 ranks x 2 copy-issuing threads oversubscribe the host
    workers = [Worker() for _ in range(nworkers)]

    def run_calls(ncalls):
        """ncalls blocking C-ABI calls in total, dealt to the workers (ctypes drops the GIL)"""
        if nworkers == 1:
            for _ in range(ncalls):
                workers[0].call()
            return
        errs = []

        def loop(w, k):
            try:
                for _ in range(k):
                    w.call()
            except Exception as e:   # noqa: BLE001
                errs.append(e)
        ths = [threading.Thread(target=loop, args=(w, ncalls // nworkers + (1 if i < ncalls % nworkers else 0))) for i, w in enumerate(workers)]
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        if errs:
            raise errs[0]

    run_calls(3 * nworkers)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    run_calls(args.steps * S)                # a step = S calls of B frames: the same S*B frames per step as the device-resident arm
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    # host<->device link speed on this box (the e2e ceiling: h2d bytes per step / link rate), pinned buffers, device timer; measured alone
    # on rank 0's turn and with ALL ranks copying at once (the ceiling the multi-GPU e2e number lives under)
    probe_d = torch.empty(64 << 20, dtype=torch.uint8, device="cuda")
    probe_h = orbx.HostBuffer((64 << 20,), np.uint8)
    probe_h.array[...] = 1
    rt = C.CDLL("libcudart.so.12")

    def probe(kind):
        dst, src = (C.c_void_p(probe_d.data_ptr()), probe_h.ptr) if kind == 1 else (probe_h.ptr, C.c_void_p(probe_d.data_ptr()))
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        rt.cudaMemcpyAsync(dst, src, C.c_size_t(64 << 20), C.c_int(kind), C.c_void_p(sh))
        torch.cuda.synchronize()
        p0.record(stream)
        for _ in range(8):
            rt.cudaMemcpyAsync(dst, src, C.c_size_t(64 << 20), C.c_int(kind), C.c_void_p(sh))
        p1.record(stream)
        torch.cuda.synchronize()
        return 8 * (64 << 20) / (p0.elapsed_time(p1) / 1e3) / 1e9

    def probe_bidir(down_share):
        """uploads and downloads at once, in the byte ratio of the workload (the e2e pipeline keeps both directions busy): GB/s of uploads"""
        s2 = torch.cuda.Stream()
        up_b, dn_b = 64 << 20, max(1 << 20, int((64 << 20) * down_share) & ~4095)
        hd = orbx.HostBuffer((dn_b,), np.uint8)
        dd = torch.empty(dn_b, dtype=torch.uint8, device="cuda")
        p0, p1, p2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        p0.record(stream)
        s2.wait_event(p0)
        for _ in range(8):
            rt.cudaMemcpyAsync(C.c_void_p(probe_d.data_ptr()), probe_h.ptr, C.c_size_t(up_b), C.c_int(1), C.c_void_p(sh))
            rt.cudaMemcpyAsync(hd.ptr, C.c_void_p(dd.data_ptr()), C.c_size_t(dn_b), C.c_int(2), C.c_void_p(s2.cuda_stream))
        p1.record(stream)
        p2.record(s2)
        torch.cuda.synchronize()
        ms = max(p0.elapsed_time(p1), p0.elapsed_time(p2))
        hd.close()
        return 8 * up_b / (ms / 1e3) / 1e9

    link = {}
    for r in range(world):              # one rank at a time
        if world > 1:
            dist.barrier()
        if r == rank:
            link = {"h2d": probe(1), "d2h": probe(2)}
    link_conc = None
    if world > 1:
        dist.barrier()
        hc = probe(1)
        dist.barrier()
        dc = probe(2)
        dist.barrier()
        bd = probe_bidir((((2 if right_records else 1) * cap * 60 + cap * 8 + 8)) / float(2 * W * H))
        t = torch.tensor([hc, dc, link["h2d"], link["d2h"], bd], device="cuda", dtype=torch.float64)
        tsum = t.clone()
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        tmin = t.clone()
        dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
        link_conc = {"h2d_sum_gbs": float(tsum[0]), "d2h_sum_gbs": float(tsum[1]), "h2d_min_rank_gbs": float(tmin[0]),
                     "h2d_alone_sum_gbs": float(tsum[2]), "d2h_alone_sum_gbs": float(tsum[3]),
                     "h2d_sum_gbs_with_downloads": float(tsum[4]),
                     "note": "all ranks copying at once; *_with_downloads = uploads while every rank also downloads in the workload's byte ratio"}
    del probe_d
    probe_h.close()
    for wi, w in enumerate(workers):   # every worker's last call must equal the device-resident results, all of them
        okc = np.array_equal(w.ocL.array, countsL) and np.array_equal(w.ocR.array, countsR)
        oku = all(w.ouR.array[f, :countsL[f]].tobytes() == uR[f, :countsL[f]].tobytes() and
                  w.odp.array[f, :countsL[f]].tobytes() == depth[f, :countsL[f]].tobytes() for f in range(B))
        if not (okc and oku):
            badc = np.nonzero(w.ocL.array != countsL)[0]
            raise AssertionError("e2e and device-resident paths disagree: worker %d, counts differ at frames %s, uRight/depth equal: %s"
                                 % (wi, badc[:8].tolist(), oku))
    h2d_call = 2 * B * W * H
    d2h_call = int(B * ((2 if right_records else 1) * cap * 60 + cap * 8 + 8))
    e2e = {"value": world * B * S * args.steps / e2e_s, "unit": "frames/s", "h2d_bytes_per_step": S * h2d_call, "d2h_bytes_per_step": S * d2h_call,
           "timer": "host wall clock around all calls", "timed_region_s": e2e_s,
           "api": "orbx_stereo_frames_host (blocking call of %d frames; a step = %d calls; %d host thread(s), each with its own handles, "
                  "take the calls in turn)" % (B, S, nworkers),
           "host_threads": nworkers, "chunks_per_call": args.e2e_chunks, "right_image_records_downloaded": right_records,
           "verified": "counts, uRight and depth of every frame of every worker's last call equal the device-resident results",
           "link_gbs_measured": {k: round(v, 2) for k, v in link.items()},
           "link_bound_frames_per_s": world * B / (h2d_call / (link["h2d"] * 1e9)),
           "link_concurrent": link_conc,
           "link_bound_frames_per_s_concurrent": (link_conc["h2d_sum_gbs_with_downloads"] * 1e9 / (2 * W * H)) if link_conc else None}

    # informational: mono extraction only at the shape the north-star target is quoted on (752x480, nFeatures 1000), device-resident and
    # end to end through orbx_extract_batch_host
    gm = orbx.ORBextractor(1000, max_batch=B, device=local)
    for _ in range(3):
        gm.extract_batch_device(devL[0].data_ptr(), B, W, H, W, W * H, (0, 0), sh)
    m0, m1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    m0.record(stream)
    nmono = max(args.steps, 10) * 4
    for k in range(nmono):
        gm.extract_batch_device(devL[k % POOL].data_ptr(), B, W, H, W, W * H, (0, 0), sh)
    m1.record(stream)
    torch.cuda.synchronize()
    mono_fps = B * nmono / (m0.elapsed_time(m1) / 1e3)
    gm.extract_batch_device(devL[0].data_ptr(), B, W, H, W, W * H, (0, 0), sh)
    torch.cuda.synchronize()
    mcount, _, mk, _ = gm.results_host(B)
    class MonoWorker:
        """one host thread with its own extractor handle and output buffers (same scheme as the stereo e2e workers above)"""

        def __init__(self):
            self.ex = orbx.ORBextractor(1000, max_batch=nb, device=local)
            self.cap = self.ex.max_keypoints(W, H)
            self.ok, self.od = hb((B, self.cap, 28), np.uint8), hb((B, self.cap, 32), np.uint8)
            self.oc, self.om = hb((B,), np.int32), hb((B,), np.int32)

        def call(self):
            rc = orbx.lib().orbx_extract_batch_host(self.ex._h, vp(hinL), B, W, H, W, C.c_size_t(W * H), 0, 0, vp(self.ok), vp(self.od), self.cap,
                                                    vp(self.oc), vp(self.om))
            if rc != 0:
                raise RuntimeError("orbx_extract_batch_host failed: %d %s" % (rc, orbx.lib().orbx_last_error()))

    mworkers = [MonoWorker() for _ in range(nworkers)]
    mcap = mworkers[0].cap

    def mono_calls(ncalls):
        if nworkers == 1:
            for _ in range(ncalls):
                mworkers[0].call()
            return
        errs = []

        def loop(w, k):
            try:
                for _ in range(k):
                    w.call()
            except Exception as e:   # noqa: BLE001
                errs.append(e)
        ths = [threading.Thread(target=loop, args=(w, ncalls // nworkers + (1 if i < ncalls % nworkers else 0))) for i, w in enumerate(mworkers)]
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        if errs:
            raise errs[0]

    mono_calls(3 * nworkers)
    if world > 1:
        dist.barrier()
    nmono_e2e = 8 * nmono                   # ~1 s of calls at the default --steps
    t0 = time.perf_counter()
    mono_calls(nmono_e2e)
    mono_e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([mono_e2e_s], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        mono_e2e_s = float(t.item())
    for w_ in mworkers:
        if not np.array_equal(w_.oc.array, mcount):
            raise AssertionError("mono e2e and device-resident keypoint counts disagree")
    mono = {"workload": "euroc_mono_752x480_nfeat1000_extract (configs[0] shape, the north-star target's shape)",
            "value": world * mono_fps if world == 1 else None, "frames_per_s_per_gpu": mono_fps, "unit": "frames/s",
            "e2e": {"value": world * B * nmono_e2e / mono_e2e_s, "unit": "frames/s",
                    "api": "orbx_extract_batch_host (blocking call of %d frames), %d host thread(s) per rank, each with its own handle" % (B, nworkers),
                    "h2d_bytes_per_call": B * W * H, "d2h_bytes_per_call": int(B * (mcap * 60 + 8)), "timed_region_s": mono_e2e_s},
            "keypoints_per_image": float(mcount.mean())}
    for o in [x for w_ in mworkers for x in (w_.ex, w_.ok, w_.od, w_.oc, w_.om)]:
        o.close()

    allpairs = None
    if not args.no_allpairs:
        allpairs = allpairs_block(orbx, torch, dist, mt, rank, world, local, args.allpairs_kf, 1000, args.allpairs_chunks)

    extras = {"mono_extract_752x480_nfeat1000_frames_per_s_per_gpu": mono_fps, "mono_1000": mono}
    if rank == 0 and not args.no_extras:
        extras.update(extra_configs(orbx, synth, torch, stream, local, max(3, min(args.steps, 10))))
    if rank == 0:
        cores = os.cpu_count() or 1
        ncpu = max(cores, min(2 * cores, 128))
        fps1, _, _, kind = cpu_stereo_sample(synth, 4, 2)
        secN, passes = 0.0, 0
        while secN < (0.0 if args.skip_cpu else 10.0) and passes < 40:   # bounded sample: ~10 s of CPU work on all host threads
            _, s_, _, kind = cpu_stereo_sample(synth, ncpu, cores, seed=2000 + passes)
            secN += s_
            passes += 1
        ncpu *= passes
        fpsN = ncpu / secN if secN > 0 else 0.0
        cpu = {"value": fpsN, "unit": "frames/s", "cores": cores, "kind": kind,
               "sample": "%d stereo frames of the workload, frame-parallel with 2 threads per frame, %d threads (%.1f s); one frame at a time "
                         "(2 threads): %.1f frames/s" % (ncpu, cores, secN, fps1), "two_thread_value": fps1,
               "note": CPU_NOTE if kind == "reference" else "oracle port (oracle/_ref not built)"}
        if not args.skip_cpu:
            cv = cv2_primitive_times(synth)
            cpu["cv2_primitives"] = cv
            if cv:   # what a SIMD OpenCV build could at best reach on this host if only those primitives cost anything
                cpu["cv2_primitive_floor_frames_per_s_all_cores"] = cores / (cv["ms_per_stereo_frame_lower_bound"] / 1e3)
        launches = (2 * exL.launches_per_batch() + STEREO_KERNELS) * S * args.steps * world
        line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": nwarm,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
                "data": "synthetic", "config": static_config(),
                "run": {"frames_per_step_per_gpu": B * S, "frames_per_launch": B, "launches_per_step": S, "timed_region_s": ms / 1e3,
                        "frames_timed_per_gpu": B * S * args.steps, "images_per_s": 2 * value, "keypoints_per_image": kp_per_image,
                        "stereo_matches_per_frame": matched, "device_pool_subbatches": POOL,
                        "working_set_mb_per_launch": 2 * B * (W * H + 2 * LEVEL_PIXELS) / 1e6, "cpu_affinity": affinity},
                "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roof, "cpu_baseline": cpu, "allpairs": allpairs,
                "extra": extras}
        _emit(line)
    for w in workers:
        for o in w.buffers():
            o.close()
    for o in [exL, exR, mt, gm, hinL, hinR]:
        o.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="orbx", choices=["orbx", "reference"])
    ap.add_argument("--batch", type=int, default=256, help="stereo frames per launch per GPU")
    ap.add_argument("--substeps", type=int, default=40, help="launches per step: a step is substeps*batch stereo frames per GPU")
    ap.add_argument("--e2e-chunks", type=int, default=4, help="chunks per orbx_stereo_frames_host call (copy/compute overlap)")
    ap.add_argument("--e2e-workers", type=int, default=0, help="host threads issuing the blocking e2e call on their own handles (0 = 2, or 1 from 4 ranks up)")
    ap.add_argument("--e2e-no-right", action="store_true", help="leave the right image's keypoints/descriptors on the device (rectified stereo needs only uRight/depth)")
    ap.add_argument("--allpairs-kf", type=int, default=2000, help="keyframes of the all-pairs sweep (x 1000 descriptors), sharded over the ranks")
    ap.add_argument("--allpairs-chunks", type=int, default=8)
    ap.add_argument("--no-allpairs", action="store_true", help="skip the all-pairs keyframe 2-NN sweep (configs[4])")
    ap.add_argument("--no-extras", action="store_true", help="skip the informational runs of BASELINE.json configs 2-4")
    ap.add_argument("--skip-cpu", action="store_true", help="development only: skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_orbx(args)


def _emit(line):
    """the ONE JSON line goes to the process's original stdout"""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


# Native libraries write to file descriptor 1 behind Python's back (NCCL prints its version banner there when the raw communicator of
# the all-pairs sweep is created): the contract is ONE JSON line on stdout, so everything else is routed to stderr at fd level.
_REAL_STDOUT = 1
if __name__ == "__main__":
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    main()
