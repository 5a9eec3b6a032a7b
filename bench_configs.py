#!/usr/bin/env python
"""Secondary bench: the other BASELINE.json configurations through the same C-ABI path as bench.py (bench.py itself times config 2, the
configuration the metric is quoted on; these lines go to profiles/, they are not the driver's bench line).

  python bench_configs.py --config 3   5 x 720x540, 2000 features per camera, semantic masks applied in the triangulation kernel
  python bench_configs.py --config 4   8 x 1920x1200, 4000 features per camera (28 camera pairs)

`value`: images (and masks) resident in HBM, results left on the device (mcs_rig_process_device); `e2e`: pinned host images and masks
through mcs_rig_process with H2D / D2H inside the timed region.  CUDA events on the rig's stream, >= 3 warm-up steps, per-step working
set far above the 126 MB L2.  Algorithmic bytes per rig frame: SURVEY.md 8(d) (bench.algorithmic_bytes with this configuration).
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    3: dict(w=720, h=540, ncam=5, nfeat=2000, masks=True, frames=256,
            name="config3: 5-camera 720x540 rig, 2000 features per camera, semantic masks applied in-kernel"),
    4: dict(w=1920, h=1200, ncam=8, nfeat=4000, masks=False, frames=32,
            name="config4: 8-camera 1920x1200 rig, 4000 features per camera, 8 levels (28 camera pairs)"),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, default=3, choices=sorted(CONFIGS))
    ap.add_argument("--frames", type=int, default=0)
    ap.add_argument("--distinct", type=int, default=0, help="ray-cast scenes (default 2; 1 for config 4)")
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    a = ap.parse_args()
    import bench
    import mcslam_b200 as M
    from mcslam_b200 import binding as B
    c = CONFIGS[a.config]
    w, h, ncam, nfeat = c["w"], c["h"], c["ncam"], c["nfeat"]
    F = a.frames or c["frames"]
    a.distinct = a.distinct or (1 if a.config == 4 else 2)
    calib = M.synth.make_rig(ncam, w, h)
    base = [M.synth.make_rig_frame(calib, w, h, frame_index=i) for i in range(a.distinct)]
    frames = np.empty((F, ncam, h, w), np.uint8)
    for f in range(F):
        frames[f] = np.roll(base[f % a.distinct], 8 * (f // a.distinct), axis=2)
    masks = None
    if c["masks"]:
        mb = [M.synth.make_masks(ncam, w, h, frame_index=i) for i in range(a.distinct)]
        masks = np.empty((F, ncam, h, w), np.float32)
        for f in range(F):
            masks[f] = np.roll(mb[f % a.distinct], 8 * (f // a.distinct), axis=2)
    rig = M.RigFrames(calib, w, h, nfeatures=nfeat, max_frames=F)
    pin = B.pinned_empty(frames.shape, np.uint8)
    pin.arr[...] = frames
    d_images = B.device_malloc(0, frames.nbytes)
    B.device_upload(0, d_images, frames)
    pin_m, d_masks = None, None
    if masks is not None:
        pin_m = B.pinned_empty(masks.shape, np.float32)
        pin_m.arr[...] = masks
        d_masks = B.device_malloc(0, masks.nbytes)
        B.device_upload(0, d_masks, masks)

    def timed(fn):
        for _ in range(max(a.warmup, 3)):
            fn()
        rig.synchronize()
        rig.timer_start()
        for _ in range(a.steps):
            fn()
        return rig.timer_stop() / a.steps

    ms_dev = timed(lambda: rig.process_device(d_images, F, d_masks))
    ms_e2e = timed(lambda: rig.process(pin.arr, masks=None if pin_m is None else pin_m.arr))
    Fp = min(F, 64 if a.config == 3 else 8)
    rig.set_profiling(True)
    acc = dict.fromkeys(bench.STAGES, 0.0)
    for _ in range(a.steps):
        rig.process_device(d_images, Fp, d_masks)
        for k, v in rig.stage_times().items():
            acc[k] += v
    rig.set_profiling(False)
    rig.process(pin.arr, masks=None if pin_m is None else pin_m.arr)
    kc, n_rows, n_acc = rig.counts(0)
    total_bytes, _ = bench.algorithmic_bytes(ncam, w, h, nfeat)
    if masks is not None:
        total_bytes += 4 * ncam * nfeat  # one mask lookup per keypoint
    peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    value = F / (ms_dev / 1e3)
    print(json.dumps({
        "metric": "rig-frames/sec (ORB extract+intra-match+triangulate)", "value": value, "unit": "rig-frames/s", "n_gpus": 1, "steps": a.steps,
        "warmup": max(a.warmup, 3), "ms_per_step": ms_dev, "dtype": "u8", "data": "synthetic",
        "config": {"workload": c["name"], "frames_per_step_per_gpu": F, "distinct_scenes": a.distinct},
        "e2e": {"value": F / (ms_e2e / 1e3), "unit": "rig-frames/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": int(frames.nbytes + (masks.nbytes if masks is not None else 0))},
        "roofline": {"bound": "hbm", "path": {"bytes_per_rig_frame": int(total_bytes), "achieved": total_bytes * value / 1e9,
                                              "frac": total_bytes * value / 1e9 / peak, "unit": "GB/s", "peak": peak}},
        "stage_ms_per_launch_sequence": {"frames": Fp, **{k: round(v / a.steps, 4) for k, v in acc.items()}},
        "counts_frame0": {"keypoints": kc.tolist(), "intramatch_rows": int(n_rows), "landmarks": int(n_acc)},
    }), flush=True)


if __name__ == "__main__":
    main()
