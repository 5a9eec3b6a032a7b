#!/usr/bin/env python
"""bench.py — rig-frames/s of the MC-SLAM front-end hot path (ORB extract + intra-rig match + triangulate) on B200.

Contract (driver): `python bench.py --gpus N --steps K --warmup W` prints ONE JSON line on rank 0.  For N > 1 it is launched
under torchrun (one rank per GPU); frames are sharded by rig frame, no collective on the data path (weak scaling: every
rank processes its own batch of `--frames` rig frames per step, generated from its own scene seeds).

  value     : rig-frames/s, inputs resident in HBM, results left on the device (mcs_rig_process_device)
  e2e       : rig-frames/s through the host-buffer C-ABI call (mcs_rig_process): pinned H2D of the images (and masks) and D2H of
              keypoints / descriptors / IntraMatch rows / landmarks inside the timed region
  roofline  : dominant stage, algorithmic bytes (DESIGN.md "Algorithmic bytes") / its CUDA-event time, vs MEASURED_PEAKS.json;
              roofline.issue = warp instructions (profiles/issue.json, from the committed ncu captures) against the SMs' issue rate
  latency   : N = 1 only: per-call latency of batches of 1 / 8 / 64 rig frames (stream path, CUDA-graph replay, host buffers)
  cpu_baseline : the reference's own code compiled here (oracle/_ref, kind "reference") when it is present and the rig has <= 5
              cameras, else the CPU restatement oracle/liboracle.so (kind "port"); bounded sample, rank 0, N = 1
`--impl reference` times that CPU path alone with all host threads.  `--config 2|3|4|noise` selects the BASELINE.json configuration
(default 2, the one the metric is quoted on; `noise` = config 2 on uniform-noise images, the FAST / octree worst case).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

WORKLOAD = "config2: 5-camera 720x540 rig, ORB(1000 features, 8 levels, 1.2, FAST 20/7) extract + 10-pair Hamming intra-match + N-view triangulation"
W, H, NCAM, NFEAT = 720, 540, 5, 1000
CONFIGS = {
    "2": dict(w=720, h=540, ncam=5, nfeat=1000, masks=False, frames=512, distinct=4, noise=False, workload=WORKLOAD),
    "3": dict(w=720, h=540, ncam=5, nfeat=2000, masks=True, frames=256, distinct=2, noise=False,
              workload="config3: 5-camera 720x540 rig, 2000 features per camera, semantic masks applied in the triangulation kernel"),
    "4": dict(w=1920, h=1200, ncam=8, nfeat=4000, masks=False, frames=32, distinct=1, noise=False,
              workload="config4: 8-camera 1920x1200 rig, 4000 features per camera, 8 levels (28 camera pairs)"),
    "noise": dict(w=720, h=540, ncam=5, nfeat=1000, masks=False, frames=256, distinct=4, noise=True,
                  workload="config2 geometry on uniform-noise images (worst case for FAST candidates and the octree)"),
}
STAGES = ["upload", "pyramid", "fast", "octree", "blur", "describe", "match", "merge", "triangulate", "download"]


def level_sizes(w, h, nlevels=8, scale=1.2):
    """Level sizes as the reference constructor + ComputePyramid derive them (f32 running product, cvRound)."""
    sf = [np.float32(1.0)]
    for _ in range(1, nlevels):
        sf.append(np.float32(np.float64(sf[-1]) * np.float64(np.float32(scale))))
    out = []
    for s in sf:
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))))
    return out


def algorithmic_bytes(ncam=NCAM, w=W, h=H, nfeat=NFEAT):
    """SURVEY.md 8(d) two-pass pixel model, per rig frame; also the per-stage split used for the dominant-kernel roofline."""
    px = [a * b for a, b in level_sizes(w, h)]
    allpx, src = sum(px), px[0]
    per_cam = {
        "pyramid": sum(px[:-1]) + sum(px[1:]),           # each level read once to make the next, levels 1..7 written once
        "fast": allpx + 4 * 10 * nfeat // 4,             # every level read once; candidate records (4 B, ~2.5 per kept feature)
        "octree": 4 * 10 * nfeat // 4 + 4 * nfeat,       # candidate records in, selected records out
        "blur": 2 * allpx,                               # every level read and written once
        "describe": (1024 + 32 + 28 + 4) * nfeat,        # 32x32 patch + selected record in, descriptor + keypoint out (blur fused)
    }
    frame = {k: v * ncam for k, v in per_cam.items()}
    frame["match"] = ncam * (ncam - 1) * nfeat * 32 + (ncam * (ncam - 1) // 2) * nfeat * 8
    frame["merge"] = (ncam * (ncam - 1) // 2) * nfeat * 4 + 3000 * ncam * 4
    frame["triangulate"] = 3000 * (4 * ncam + 24)
    extract = ncam * (src + (allpx - src) + allpx + 60 * nfeat)
    total = extract + frame["match"] + frame["triangulate"]
    return total, frame


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(device), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 7:
                continue
            try:
                sm.append(float(c[0])); mx.append(float(c[1])); pw.append(float(c[2]))
            except ValueError:
                continue
            for n, v in zip(names, c[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        self.f.close()
        os.unlink(self.f.name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)), "samples": len(sm),
                "reasons": sorted(reasons)}


def make_workload(n_frames, distinct, rank, cfg=None, world=1):
    """This rank's share of ONE stream of world * n_frames rig frames: [n_frames, ncam, h, w] u8 (+ f32 masks for config 3) and the
    global frame index of every local frame.  Rig frame g of the stream belongs to rank g mod world (mcslam_b200.sharding).
    Frame g shows scene g mod (distinct * world) — rendered by the synthetic scene generator, seed = 1234 + scene —
    cyclically shifted by 8 * (g div (distinct * world)) columns (all cameras alike, so the cross-camera geometry is kept): distinct
    pixel data per frame without minutes of ray casting; every rank renders `distinct` scenes of its own."""
    import mcslam_b200 as M
    from mcslam_b200 import sharding as S
    cfg = cfg or CONFIGS["2"]
    w, h, ncam = cfg["w"], cfg["h"], cfg["ncam"]
    calib = M.synth.make_rig(ncam, w, h)
    mine = S.frames_for_rank(world * n_frames, rank, world)
    period = distinct * world
    scenes, mscenes = {}, {}
    frames = np.empty((n_frames, ncam, h, w), np.uint8)
    masks = np.empty((n_frames, ncam, h, w), np.float32) if cfg["masks"] else None
    for j, g in enumerate(mine):
        sc = g % period
        if sc not in scenes:
            scenes[sc] = (np.stack([M.synth.noise_image(w, h, 977 * sc + c) for c in range(ncam)]) if cfg["noise"]
                          else M.synth.make_rig_frame(calib, w, h, frame_index=sc))
            if masks is not None:
                mscenes[sc] = M.synth.make_masks(ncam, w, h, frame_index=sc)
        frames[j] = np.roll(scenes[sc], 8 * (g // period), axis=2)
        if masks is not None:
            masks[j] = np.roll(mscenes[sc], 8 * (g // period), axis=2)
    return calib, frames, masks, mine


class CpuArm:
    """The CPU implementation of the path for one configuration: oracle/_ref (the reference's own ORBextractor.cpp and MultiCameraFrame /
    FrontEnd method bodies compiled here; IntraMatch holds 5 cameras) when present, else the restatement oracle/liboracle.so."""

    def __init__(self, cfg, calib, want=1):
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import oracle_lib as O
        self.cfg, self.calib, self.O = cfg, calib, O
        self.kind = "port"
        self.R = None
        try:
            import ref_lib as R
            if R.available() and cfg["ncam"] <= 5 and not os.environ.get("MCS_BENCH_CPU_PORT"):
                self.R, self.kind = R, "reference"
        except Exception:
            pass
        self.workers = [self._make() for _ in range(want)]

    def _make(self):
        c = self.cfg
        if self.R is not None:
            return self.R.Rig(self.calib, c["w"], c["h"], nfeatures=c["nfeat"])
        return [self.O.Extractor(c["nfeat"]) for _ in range(c["ncam"])]

    def run(self, i, images, do_match=True, do_triangulate=True):
        """One rig frame on worker i, one thread per camera for the extraction as the reference does (MultiCameraFrame.cpp:212-227)."""
        wk = self.workers[i]
        if self.R is not None:
            return wk.process(images, do_match, do_triangulate)
        return self.O.process_rig(wk, images, self.calib, threads=self.cfg["ncam"], do_match=do_match, do_triangulate=do_triangulate)

    def stage_ms(self, frames, n=6):
        """Per-stage time of one rig frame (the reference logs the same three, FrontEnd.cpp:992-1014): extraction alone, then with the
        intra-match, then with the triangulation, on the same frame; differences of the fastest of n runs each (the stages are cumulative
        calls, so a difference of noisy medians could come out negative), floored at zero."""
        t = []
        for flags in ((False, False), (True, False), (True, True)):
            dt = []
            for k in range(n):
                t0 = time.perf_counter()
                self.run(0, frames[0], *flags)
                dt.append(time.perf_counter() - t0)
            t.append(1e3 * float(min(dt)))
        return {"extract": round(t[0], 2), "intra_match": round(max(t[1] - t[0], 0.0), 2), "triangulate": round(max(t[2] - t[1], 0.0), 2)}


def cv2_primitives_ms(frames, nfeat, n=2):
    """Sanity figure SURVEY.md 8(d) asks for where cv2 exists: the OpenCV primitives the reference's extractor and matcher spend their
    time in, called through cv2 on the bench's own images with the reference's call shapes — per camera 7 x resize + copyMakeBorder,
    FAST(20) per 35 px cell with the minTh retry on empty cells, 8 x GaussianBlur(7x7, 2); per rig frame BFMatcher(HAMMING).knnMatch(k=2)
    for every camera pair on random descriptors of the configured count.  No octree, no descriptor sampling, no merge, no SVD: a lower
    bound of an OpenCV-backed reference, not a replacement for the timed arm.  OpenCV's own threads.  None when cv2 is absent."""
    try:
        import cv2
    except Exception:
        return None
    ncam = frames.shape[1]
    fast_i = cv2.FastFeatureDetector_create(20, True)
    fast_m = cv2.FastFeatureDetector_create(7, True)
    rng = np.random.default_rng(0)
    descs = [rng.integers(0, 256, (nfeat, 32), dtype=np.uint8) for _ in range(ncam)]
    bf = cv2.BFMatcher(cv2.NORM_HAMMING)
    t_ext, t_match = [], []
    for k in range(n):
        t0 = time.perf_counter()
        for c in range(ncam):
            img = frames[k % len(frames), c]
            level = img
            for l, (lw, lh) in enumerate(level_sizes(img.shape[1], img.shape[0])):
                if l:
                    level = cv2.resize(level, (lw, lh), interpolation=cv2.INTER_LINEAR)
                cv2.copyMakeBorder(level, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
                W = 35
                wd, ht = lw - 32, lh - 32
                ncol, nrow = max(wd // W, 1), max(ht // W, 1)
                wc, hc = -(-wd // ncol), -(-ht // nrow)
                for i in range(nrow):
                    y0 = 16 + i * hc
                    if y0 >= lh - 16 - 3:
                        continue
                    y1 = min(y0 + hc + 6, lh - 16)
                    for j in range(ncol):
                        x0 = 16 + j * wc
                        if x0 >= lw - 16 - 6:
                            continue
                        cell = level[y0:y1, x0:min(x0 + wc + 6, lw - 16)]
                        if not fast_i.detect(cell):
                            fast_m.detect(cell)
                cv2.GaussianBlur(level, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        t1 = time.perf_counter()
        for i in range(ncam - 1):
            for j in range(i + 1, ncam):
                bf.knnMatch(descs[i], descs[j], k=2)
        t2 = time.perf_counter()
        t_ext.append(1e3 * (t1 - t0)); t_match.append(1e3 * (t2 - t1))
    return {"extract_primitives_ms": round(float(np.median(t_ext)), 2), "knn_match_ms": round(float(np.median(t_match)), 2),
            "cv2": cv2.__version__, "cv2_threads": int(cv2.getNumThreads()), "note": "primitives only, sequential over the cameras; a lower bound"}


def cpu_sample(arm, frames, seconds, pool_threads):
    """Times the CPU arm on as many frames of the workload as fit in ~`seconds`; pool_threads rig frames run concurrently, each with the
    reference's thread-per-camera fan-out."""
    from concurrent.futures import ThreadPoolExecutor
    arm.run(0, frames[0])  # warm-up (page-in, allocations)
    done = 0
    t0 = time.perf_counter()
    with ThreadPoolExecutor(pool_threads) as tp:
        while True:
            list(tp.map(lambda i: arm.run(i, frames[(done + i) % len(frames)]), range(pool_threads)))
            done += pool_threads
            if time.perf_counter() - t0 >= seconds:
                break
    dt = time.perf_counter() - t0
    return done / dt, done, dt


def run_reference(args, rank, cfg):
    """--impl reference: the CPU path on the box's host cores — the reference's own compiled code (oracle/_ref) where it exists."""
    if rank != 0:
        return
    ncores = os.cpu_count() or 1
    ncam = cfg["ncam"]
    pool = max(1, ncores // ncam)
    calib, frames, _, _ = make_workload(max(pool, 2), min(2, cfg["distinct"]), 0, cfg)
    arm = CpuArm(cfg, calib, pool)
    per_step = []
    for s_ in range(args.warmup + args.steps):
        fps, n, dt = cpu_sample(arm, frames, args.cpu_step_seconds, pool)
        if s_ >= args.warmup:
            per_step.append((fps, n, dt))
    fps = float(np.mean([p[0] for p in per_step]))
    line = {
        "impl": "reference", "metric": "rig-frames/sec (ORB extract+intra-match+triangulate)", "value": fps, "unit": "rig-frames/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * float(np.mean([p[2] for p in per_step])),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": cfg["workload"], "frames_per_step": int(np.mean([p[1] for p in per_step])), "note": "CPU only; bounded sample per step"},
        "cpu_baseline": {"value": fps, "unit": "rig-frames/s", "cores": min(ncores, pool * ncam), "kind": arm.kind,
                         "sample": "%d concurrent rig frames x %d camera threads, ~%.0f s per step" % (pool, ncam, args.cpu_step_seconds),
                         "stage_ms_one_frame": arm.stage_ms(frames)},
        "e2e": {"value": fps, "unit": "rig-frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def note(rank, msg):
    """Progress line on stderr (stdout carries only the JSON line): which phase a rank is in when a run is cut short."""
    print("[bench rank %d] %s" % (rank, msg), file=sys.stderr, flush=True)


def dist_setup(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def run_gpu(args, rank, world, local, cfg):
    import mcslam_b200 as M
    from mcslam_b200 import binding as B
    if M.lib().mcs_device_count() < 1:
        raise SystemExit("bench.py: no CUDA device — the product has no CPU path (use --impl reference for the CPU arm)")
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    # one process per GPU: keep this rank's host thread and its pinned staging buffers on the GPU's own NUMA node
    cpus_bound = 0 if os.environ.get("MCS_NO_AFFINITY") else M.lib().mcs_bind_host_thread(local)
    F = args.frames
    w, h, ncam, nfeat = cfg["w"], cfg["h"], cfg["ncam"], cfg["nfeat"]
    note(rank, "rendering the workload")
    calib, frames, masks, mine = make_workload(F, args.distinct, rank, cfg, world)
    note(rank, "workload ready, creating the rig handle")
    rig = M.RigFrames(calib, w, h, nfeatures=nfeat, max_frames=F, device=local)
    # pinned host staging (the e2e leg copies from here every step) + a resident device copy (the `value` leg)
    pin = B.pinned_empty(frames.shape, np.uint8)
    pin.arr[...] = frames
    d_images = B.device_malloc(local, frames.nbytes)
    B.device_upload(local, d_images, frames)
    pin_m = pin_u8 = None
    d_masks = None
    if masks is not None:
        pin_m = B.pinned_empty(masks.shape, np.float32)
        pin_m.arr[...] = masks
        pin_u8 = B.pinned_empty(masks.shape, np.uint8)
        pin_u8.arr[...] = np.where(masks.astype(np.float64) < 0.7, 0, 255)
        d_masks = B.device_malloc(local, masks.nbytes)
        B.device_upload(local, d_masks, masks)

    def barrier():
        rig.synchronize()
        if dist is not None:
            dist.barrier()

    def max_over_ranks(ms):
        if dist is None:
            return ms
        import torch
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed(fn, steps):
        barrier()
        rig.timer_start()
        for _ in range(steps):
            fn()
        ms = rig.timer_stop()
        barrier()
        return max_over_ranks(ms)

    dev_step = lambda: rig.process_device(d_images, F, d_masks)                            # noqa: E731
    host_step = lambda: rig.process(pin.arr, masks=None if pin_m is None else pin_m.arr)   # noqa: E731
    note(rank, "warm-up")
    for _ in range(max(args.warmup, 3)):
        dev_step()
    rig.synchronize()
    note(rank, "timing the device-resident leg")
    launches0 = rig.kernel_launches
    clocks = ClockSampler(local)
    ms_dev = timed(dev_step, args.steps)
    launches = rig.kernel_launches - launches0
    note(rank, "timing the host-buffer leg")
    for _ in range(max(args.warmup, 3)):
        host_step()
    ms_e2e = timed(host_step, args.steps)
    ms_e2e_u8 = None
    if pin_u8 is not None:  # the same call with the mask decision as one byte per pixel (mcs_rig_process_masks_u8)
        u8_step = lambda: rig.process_masks_u8(pin.arr, pin_u8.arr)                        # noqa: E731
        for _ in range(max(args.warmup, 3)):
            u8_step()
        ms_e2e_u8 = timed(u8_step, args.steps)
    clk = clocks.stop()
    # per-stage device times (CUDA events between the stages on the rig's stream) with profiling on: one unchunked launch
    # sequence over `Fp` frames per step, i.e. exactly what one pipeline chunk launches (and what the ncu captures profile)
    Fp = min(F, 64 if cfg["ncam"] * cfg["w"] * cfg["h"] < 8e6 else 8)
    rig.set_profiling(True)
    acc = dict.fromkeys(STAGES, 0.0)
    for _ in range(args.steps):
        rig.process_device(d_images, Fp, d_masks)
        for k, v in rig.stage_times().items():
            acc[k] += v
    rig.set_profiling(False)
    stage_ms = {k: v / args.steps for k, v in acc.items()}
    latency = None
    if world == 1 and not args.no_latency:
        latency = latency_sweep(rig, d_images, d_masks, pin, pin_m, F, args.latency_iters)
    host_step()  # leave the results of a full (chunked, host-buffer) step in the handle for the parity block
    kc, n_rows, n_acc = rig.counts(0)
    d2h = int(rig.last_d2h_bytes)  # counted by the library from the copies of that step (rows / landmarks sized by counts)
    # what a multi-GPU front-end hands on: per-frame counts gathered to rank 0 and ordered by global frame index (the features themselves
    # stay with the rank that owns the frame); after the timed region
    gathered = None
    note(rank, "gathering the per-frame counts")
    try:
        from mcslam_b200 import sharding as S
        local = {}
        for j, g in enumerate(mine):
            c = rig.counts(j)
            local[g] = (int(c[0].sum()), int(c[1]), int(c[2]))
        allf = S.gather_by_frame(local, world * F, dist)
        if allf is not None:
            gathered = {"frames": len(allf), "keypoints": sum(a[0] for a in allf), "intramatch_rows": sum(a[1] for a in allf),
                        "landmarks": sum(a[2] for a in allf), "order": "global frame index, frame g from rank g mod n_gpus"}
    except Exception as e:  # reported, does not void the timing
        gathered = {"error": str(e)[:200]}
    note(rank, "done")
    h2d = frames.nbytes + (masks.nbytes if masks is not None else 0)
    return dict(rig=rig, calib=calib, frames=frames, masks=masks, ms_dev=ms_dev, ms_e2e=ms_e2e, ms_e2e_u8=ms_e2e_u8, launches=launches, clocks=clk,
                stage_ms=stage_ms, cpus_bound=cpus_bound, counts=(kc.tolist(), n_rows, n_acc), d2h=d2h, h2d=h2d, F=F, Fp=Fp, dist=dist, latency=latency,
                gathered=gathered)


def latency_sweep(rig, d_images, d_masks, pin, pin_m, F, iters):
    """SURVEY.md 8(d) config 5 / FrontEnd::processFrame is called once per rig frame: host-observed latency (call -> results complete on
    the device, resp. in host memory) of batches of 1, 8 and 64 rig frames.  `device`: mcs_rig_process_device on the streams;
    `device_graph`: the same call replayed from a CUDA graph (mcs_rig_set_graph); `host`: mcs_rig_process from pinned buffers with the
    results copied back."""
    out = {}
    for b in (1, 8, 64):
        if b > F:
            continue
        n = max(20, iters // (1 if b < 64 else 4))
        row = {}
        for name in ("device", "device_graph", "host"):
            rig.set_graph(name == "device_graph")
            if name == "host":
                fn = lambda: rig.process(pin.arr[:b], masks=None if pin_m is None else pin_m.arr[:b])  # noqa: E731
            else:
                fn = lambda: (rig.process_device(d_images, b, d_masks), rig.synchronize())              # noqa: E731
            try:
                for _ in range(5):
                    fn()
            except RuntimeError as e:  # a mode that cannot run is reported, it does not void the line
                row[name] = {"error": str(e)[:200]}
                continue
            t = []
            for _ in range(n):
                t0 = time.perf_counter()
                fn()
                t.append(1e3 * (time.perf_counter() - t0))
            t = np.sort(np.array(t))
            row[name] = {"p50_ms": round(float(np.median(t)), 4), "p99_ms": round(float(t[min(len(t) - 1, int(0.99 * len(t)))]), 4),
                         "rig_frames_per_s": round(b / (float(np.median(t)) / 1e3), 1)}
        rig.set_graph(False)
        if b <= 8:  # where a small call's time goes: CUDA events between the stages of one unchunked call
            rig.set_profiling(True)
            acc = dict.fromkeys(STAGES, 0.0)
            for _ in range(20):
                rig.process_device(d_images, b, d_masks)
                for k, v in rig.stage_times().items():
                    acc[k] += v / 20
            rig.set_profiling(False)
            row["stage_ms"] = {k: round(v, 4) for k, v in acc.items() if k not in ("upload", "download", "blur")}
        out["batch_%d" % b] = row
    out["iters"] = iters
    out["clock"] = "host perf_counter around the call + stream synchronize"
    return out


def parity_report(rig, calib, frames, masks, nfeat):
    """Mismatch counts against the CPU oracle (keypoints, descriptor bits, IntraMatch rows, 3-D points) for rig frames taken from
    the first, a middle and the last pipeline chunk of the last timed step."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_lib as O
    ex = O.Extractor(nfeat)
    F = len(frames)
    picks = sorted({0, F // 2 + 1 if F > 2 else 0, F - 1})
    bad_kp = bad_bits = bad_rows = bad_status = 0
    max_rel = 0.0
    for f in picks:
        K, D = [], []
        for c in range(rig.C):
            k, d = rig.features(f, c)
            _, k2, d2 = ex(frames[f, c])
            K.append(k2); D.append(d2)
            if len(k) != len(k2):
                bad_kp += abs(len(k) - len(k2)) + 1
                continue
            bad_kp += int((k != k2).sum())
            bad_bits += int(np.unpackbits(d ^ d2).sum())
        rows, _ = O.intra_matches(D, K, calib, ex.sigma2)
        got = rig.intra_matches(f)
        if got.shape != rows.shape:
            bad_rows += abs(len(got) - len(rows)) + 1
            continue
        bad_rows += int((got != rows).any(axis=1).sum())
        tri = O.triangulate_rows(calib, K, D, rows, masks=None if masks is None else masks[f])
        lm = rig.landmarks(f)
        bad_status += int((lm["status"] != tri["status"]).sum())
        ok = tri["nviews"] >= 2
        rel = np.abs(lm["xyz"][ok] - tri["xyz"][ok]) / np.maximum(np.abs(tri["xyz"][ok]), 1e-9)
        if rel.size:
            max_rel = max(max_rel, float(np.nanmax(rel)))
    return {"frames": picks, "mismatched_keypoints": bad_kp, "mismatched_descriptor_bits": bad_bits, "mismatched_intramatch_rows": bad_rows,
            "mismatched_depth_gate": bad_status, "xyz_max_rel_err": max_rel, "xyz_tolerance": 1e-6}


def issue_roofline(stage_ms, Fp, amortised_ms_per_seq, clocks, dom):
    """Instruction-issue roofline: warp instructions per launch sequence of `frames_per_sequence` rig frames, per kernel, from the committed
    ncu captures (profiles/issue.json, written by tools/issue_table.py from profiles/*_raw.csv: smsp__inst_executed.sum), against
    SMs x 4 schedulers x SM clock.  The path is integer work on bytes: this, not HBM, is the roofline that binds it."""
    path = os.path.join(ROOT, "profiles", "issue.json")
    if not os.path.exists(path):
        return None
    t = json.load(open(path))
    if t.get("frames_per_sequence") != Fp:
        return None
    mhz = clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or t.get("sm_mhz_fallback", 1965.0)
    peak = t.get("sms", 148) * 4 * mhz * 1e6  # warp instructions per second
    total = sum(k["warp_insts"] for k in t["kernels"].values())
    out = {"unit": "warp-inst/s", "peak": peak, "sm_mhz": mhz, "sms": t.get("sms", 148), "frames_per_sequence": Fp,
           "warp_insts_per_sequence": total, "per_kernel": {k: v["warp_insts"] for k, v in t["kernels"].items()},
           "path": {"achieved": total / (amortised_ms_per_seq / 1e3), "frac": total / (amortised_ms_per_seq / 1e3) / peak,
                    "ms_at_full_issue_rate": 1e3 * total / peak},
           "source": "profiles/issue.json"}
    if dom in t["kernels"] and stage_ms.get(dom):
        wi = t["kernels"][dom]["warp_insts"]
        out["kernel"] = {"name": dom, "warp_insts": wi, "achieved": wi / (stage_ms[dom] / 1e3), "frac": wi / (stage_ms[dom] / 1e3) / peak}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="2", choices=sorted(CONFIGS), help="BASELINE.json configuration (default 2: the headline metric)")
    ap.add_argument("--frames", type=int, default=0, help="rig frames per step per GPU (default: 512 for config 2)")
    ap.add_argument("--distinct", type=int, default=0, help="ray-cast scenes per rank; the batch is filled with shifted copies")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="CPU baseline sample length")
    ap.add_argument("--cpu-step-seconds", type=float, default=6.0, help="--impl reference: seconds of CPU work per step")
    ap.add_argument("--no-cpu", action="store_true", help="skip the oracle parity check and the CPU baseline (profiling runs)")
    ap.add_argument("--no-latency", action="store_true", help="skip the batch 1 / 8 / 64 latency sweep")
    ap.add_argument("--latency-iters", type=int, default=200)
    args = ap.parse_args()
    cfg = CONFIGS[args.config]
    args.frames = args.frames or cfg["frames"]
    args.distinct = args.distinct or cfg["distinct"]
    rank, world, local = dist_setup(args)
    if args.gpus > 1 and world == 1 and "RANK" not in os.environ:
        # launched bare with --gpus N: re-launch under torchrun, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus), "--master-addr", "127.0.0.1",
               "--master-port", str(29500 + os.getpid() % 1000), os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    if args.impl == "reference":
        run_reference(args, rank, cfg)
        return
    r = run_gpu(args, rank, world, local, cfg)
    F, steps = r["F"], args.steps
    w, h, ncam, nfeat = cfg["w"], cfg["h"], cfg["ncam"], cfg["nfeat"]
    total_bytes, stage_bytes = algorithmic_bytes(ncam, w, h, nfeat)
    if cfg["masks"]:
        total_bytes += 4 * ncam * nfeat  # one mask lookup per keypoint
    io_bytes = ncam * w * h + 60 * ncam * nfeat + (ncam * (ncam - 1) // 2) * nfeat * 8 + stage_bytes["triangulate"]
    value = world * F * steps / (r["ms_dev"] / 1e3)
    e2e = world * F * steps / (r["ms_e2e"] / 1e3)
    if rank != 0:
        if r["dist"] is not None:
            r["dist"].destroy_process_group()
        return
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured)"
    else:
        peak, peak_src = 6650.0, "B200_PROFILING.md fallback"
    sm, Fp = r["stage_ms"], r["Fp"]
    kernels = {k: v for k, v in sm.items() if k not in ("upload", "download")}
    dom = max(kernels, key=kernels.get)
    launches_per_step = {"pyramid": 7, "triangulate": 3}.get(dom, 1)
    achieved = stage_bytes[dom] * Fp / (sm[dom] / 1e3) / 1e9
    # DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture of the same launch shape
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath) and args.config == "2":
        t = json.load(open(tpath)).get(dom)
        if t and t.get("frames_per_launch") == Fp:
            traffic = t["dram_bytes_per_launch"]
    amortised = r["ms_dev"] / steps / (F / Fp)  # device time per Fp rig frames inside the chunked pipeline
    issue_bound = ("pyramid", "fast", "describe", "match")
    line = {
        "metric": "rig-frames/sec (ORB extract+intra-match+triangulate)", "value": value, "unit": "rig-frames/s", "n_gpus": world,
        "steps": steps, "warmup": max(args.warmup, 3), "ms_per_step": r["ms_dev"] / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": cfg["workload"], "frames_per_step_per_gpu": F, "distinct_scenes": args.distinct,
                   "chunk_frames": int(os.environ.get("MCS_CHUNK_FRAMES", "64")),
                   "l2": "per-step working set (%.0f MB of images + pyramids per GPU) exceeds the 126 MB L2" % (F * ncam * (w * h * 2.67) / 1e6),
                   "sharding": "one stream of n_gpus x frames_per_step_per_gpu rig frames, frame g on rank g mod n_gpus, no collective on the data path",
                   "host_cpus_bound_to_gpu_numa_node": int(r["cpus_bound"])},
        "e2e": {"value": e2e, "unit": "rig-frames/s", "h2d_bytes_per_step": int(r["h2d"]), "d2h_bytes_per_step": int(r["d2h"]),
                "ms_per_step": r["ms_e2e"] / steps},
        "gpu_launches": int(r["launches"]),
        "clocks": r["clocks"],
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "peak_source": peak_src, "frames_per_launch": Fp,
                     "algorithmic_bytes_per_launch": stage_bytes[dom] * Fp // launches_per_step,
                     "launch_ms": sm[dom] / launches_per_step,
                     "path": {"bytes_per_rig_frame": total_bytes, "achieved": total_bytes * value / world / 1e9,
                              "frac": total_bytes * value / world / 1e9 / peak,
                              # SURVEY.md 8(d) "also report": the inputs-and-outputs-only lower bound (source pixels + keypoints +
                              # descriptors + pair matches + points), so nobody mistakes the two-pass model for a minimum
                              "io_only_bytes_per_rig_frame": io_bytes, "io_only_frac": io_bytes * value / world / 1e9 / peak}},
        "stage_ms_per_launch_sequence": {"frames": Fp, **{k: round(v, 4) for k, v in sm.items()}},
        # how much of the latency-bound kernels (octree, merge, triangulation) the chunk pipeline hides: the amortised device time per
        # launch sequence next to the sum of the issue-bound kernels alone
        "overlap": {"frames": Fp, "amortised_ms": round(amortised, 4), "issue_bound_sum_ms": round(sum(sm[k] for k in issue_bound), 4),
                    "all_kernels_sum_ms": round(sum(kernels.values()), 4)},
        "counts_frame0": {"keypoints": r["counts"][0], "intramatch_rows": r["counts"][1], "landmarks": r["counts"][2]},
        "gathered_counts": r["gathered"],
    }
    if r["ms_e2e_u8"] is not None:
        line["e2e"]["masks_u8"] = {"value": world * F * steps / (r["ms_e2e_u8"] / 1e3), "ms_per_step": r["ms_e2e_u8"] / steps,
                                   "h2d_bytes_per_step": int(r["frames"].nbytes + r["masks"].size)}
    if args.config == "2":
        iss = issue_roofline(sm, Fp, amortised, r["clocks"], dom)
        if iss:
            line["roofline"]["issue"] = iss
    if r["latency"]:
        line["latency"] = r["latency"]
    if world == 1 and not args.no_cpu:
        line["parity"] = parity_report(r["rig"], r["calib"], r["frames"], r["masks"], nfeat)
        ncores = os.cpu_count() or 1
        arm = CpuArm(cfg, r["calib"], 1)
        fps, n, dt = cpu_sample(arm, r["frames"], args.cpu_seconds, 1)
        line["cpu_baseline"] = {"value": fps, "unit": "rig-frames/s", "cores": min(ncam, ncores), "kind": arm.kind,
                                "sample": "%d rig frames of the workload in %.1f s, reference threading (1 thread per camera)" % (n, dt),
                                "host_cores": ncores, "stage_ms_one_frame": arm.stage_ms(r["frames"]),
                                "cv2_primitives_one_frame": cv2_primitives_ms(r["frames"], nfeat)}
        if r["latency"] and "batch_1" in r["latency"]:
            line["latency"]["cpu_ms_per_rig_frame"] = round(1e3 / fps, 2)
    print(json.dumps(line), flush=True)
    if r["dist"] is not None:
        r["dist"].destroy_process_group()


if __name__ == "__main__":
    main()
