#!/usr/bin/env python
"""Copy-only probe of the host <-> device path that bench.py's `e2e` leg uses, at 1..N ranks at once — no kernels.

    python tools/pcie_probe.py                       # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 tools/pcie_probe.py

Every rank owns pinned buffers the size of one config-2 step (512 rig frames: 995 MB of pixels up, 266 MB of results down) and moves
them with the copy pattern of mcs_rig_process: chunks of 16 rig frames rotating over six streams, per chunk one H2D copy of the images
and 15 D2H copies (keypoints, descriptors, counts, rows, landmark slabs).  Three patterns are timed with CUDA events, max over ranks:
  h2d_chunks      the uploads alone           h2d_d2h_chunks   uploads and downloads together (what the e2e leg does)
  h2d_one_copy    one 995 MB copy per step (the ceiling of the link + host memory for this rank count)
Prints one JSON line on rank 0: aggregate GB/s per pattern and the rig-frames/s a copy-bound e2e leg could reach (1.944 MB up per frame).
torch is used for device memory, streams and the barrier only.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

FRAMES, CHUNK, NSLOT = 512, 16, 6
UP_PER_FRAME = 5 * 720 * 540
# the 16 result copies of one chunk, bytes per rig frame (kp_cap 1004, 5 cameras, 8 levels): download_features_async,
# download_rows_async, download_landmarks_async in mc-slam_b200/csrc/mcslam_b200.cu.  ROWS = rows per frame the row / landmark copies
# carry: the library sends the previous call's fullest frame + 25 % + 64 (about 1500 on the bench workload; MCS_PROBE_ROWS=2510 is the
# table capacity, the round-1 pattern the first committed probe run used)
ROWS = int(os.environ.get("MCS_PROBE_ROWS", "1500"))
DOWN_PER_FRAME = [20, 20, 160, 160, 5 * 1004 * 28, 5 * 1004 * 32, 4, ROWS * 20, 4, 4, ROWS * 20, ROWS, ROWS * 4, ROWS * 24, ROWS * 8, ROWS * 4]


def main():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    up_h = torch.empty(FRAMES * UP_PER_FRAME, dtype=torch.uint8).pin_memory()
    up_h.fill_(1)
    up_d = torch.empty_like(up_h, device="cuda")
    down_d = [torch.zeros(FRAMES * b, dtype=torch.uint8, device="cuda") for b in DOWN_PER_FRAME]
    down_h = [torch.empty(FRAMES * b, dtype=torch.uint8).pin_memory() for b in DOWN_PER_FRAME]
    streams = [torch.cuda.Stream() for _ in range(NSLOT)]
    main_s = torch.cuda.current_stream()

    def step(h2d, d2h, one_copy=False):
        if one_copy:
            up_d.copy_(up_h, non_blocking=True)
            return
        ev = torch.cuda.Event()
        ev.record(main_s)
        for s in streams:
            s.wait_event(ev)
        for ci, f0 in enumerate(range(0, FRAMES, CHUNK)):
            s = streams[ci % NSLOT]
            with torch.cuda.stream(s):
                if h2d:
                    a, b = f0 * UP_PER_FRAME, (f0 + CHUNK) * UP_PER_FRAME
                    up_d[a:b].copy_(up_h[a:b], non_blocking=True)
                if d2h:
                    for dh, dd, per in zip(down_h, down_d, DOWN_PER_FRAME):
                        a, b = f0 * per, (f0 + CHUNK) * per
                        dh[a:b].copy_(dd[a:b], non_blocking=True)
        for s in streams:
            e = torch.cuda.Event()
            e.record(s)
            main_s.wait_event(e)

    def timed(steps, **kw):
        for _ in range(3):
            step(**kw)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record(main_s)
        for _ in range(steps):
            step(**kw)
        t1.record(main_s)
        torch.cuda.synchronize()
        ms = torch.tensor([t0.elapsed_time(t1)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            dist.barrier()
        return float(ms.item()) / steps

    steps = 10
    up = FRAMES * UP_PER_FRAME
    down = FRAMES * sum(DOWN_PER_FRAME)
    out = {"ranks": world, "rows_per_frame_sent": ROWS, "frames_per_step": FRAMES, "chunk_frames": CHUNK, "streams": NSLOT, "h2d_bytes_per_step": up, "d2h_bytes_per_step": down}
    ms = timed(steps, h2d=True, d2h=False)
    out["h2d_chunks"] = {"ms_per_step": round(ms, 3), "aggregate_GBps": round(world * up / ms / 1e6, 1), "copy_bound_rig_frames_per_s": round(world * FRAMES / ms * 1e3)}
    ms = timed(steps, h2d=True, d2h=True)
    out["h2d_d2h_chunks"] = {"ms_per_step": round(ms, 3), "aggregate_h2d_GBps": round(world * up / ms / 1e6, 1), "aggregate_d2h_GBps": round(world * down / ms / 1e6, 1),
                             "copy_bound_rig_frames_per_s": round(world * FRAMES / ms * 1e3)}
    ms = timed(steps, h2d=True, d2h=False, one_copy=True)
    out["h2d_one_copy"] = {"ms_per_step": round(ms, 3), "aggregate_GBps": round(world * up / ms / 1e6, 1), "copy_bound_rig_frames_per_s": round(world * FRAMES / ms * 1e3)}
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    sys.exit(main())
