#!/usr/bin/env python
"""Secondary bench (not the BASELINE.json headline): the inter-frame knnMatch(k=2) primitive of
FrontEnd::findMatchesMono / findInterMatches (FrontEnd.cpp:3134, :3200) through the C ABI.

  python bench_interframe.py [--sets 1280] [--rows 1000] [--steps 20] [--warmup 3]

One step = `sets` independent (query, train) problems of `rows` x `rows` 256-bit descriptors (1280 = 256 rig frames x 5 cameras,
frame t vs t+1 per camera).  `value`: descriptors resident in HBM (mcs_knn2_device); `e2e`: host buffers through mcs_knn2 (H2D of both
descriptor sets + D2H of the int32 [rows][4] result inside the timed region).  Roofline: the kernel is INT-ALU bound (XOR + carry-save
adders + POPC + top-2 update per distance), so the bound reported is distance evaluations / s against the measured ALU issue rate.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sets", type=int, default=1280)
    ap.add_argument("--rows", type=int, default=1000)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--no-cpu", action="store_true")
    a = ap.parse_args()
    import torch
    import mcslam_b200 as M
    from mcslam_b200 import binding as B
    S, R = a.sets, a.rows
    rng = np.random.default_rng(1)
    t = rng.integers(0, 256, (S * R, 32), dtype=np.uint8)
    q = t.reshape(S, R, 32)[:, rng.permutation(R)].reshape(S * R, 32).copy()
    q ^= (rng.integers(0, 256, q.shape, dtype=np.uint8) & rng.integers(0, 256, q.shape, dtype=np.uint8) & rng.integers(0, 256, q.shape, dtype=np.uint8))
    m = M.Matcher(max_sets=S, max_query_rows=S * R, max_train_rows=S * R)
    dq, dt, do = B.device_malloc(0, q.nbytes), B.device_malloc(0, t.nbytes), B.device_malloc(0, S * R * 16)
    B.device_upload(0, dq, q)
    B.device_upload(0, dt, t)
    counts = np.full(S, R, np.int32)
    st = torch.cuda.ExternalStream(B.lib().mcs_matcher_stream(m.hd))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def timed(fn):
        for _ in range(a.warmup):
            fn()
        m.synchronize()
        e0.record(st)
        for _ in range(a.steps):
            fn()
        e1.record(st)
        e1.synchronize()
        return e0.elapsed_time(e1) / a.steps

    ms_dev = timed(lambda: m.knn2_device(dq, counts, dt, counts, do))
    out = np.zeros((S * R, 4), np.int32)
    B.device_download(0, out, do)
    qs, ts = list(q.reshape(S, R, 32)), list(t.reshape(S, R, 32))
    t0 = time.perf_counter()
    for _ in range(3):
        host = m.knn2(qs, ts)
    ms_e2e = (time.perf_counter() - t0) / 3 * 1e3
    assert all(np.array_equal(h, out[i * R:(i + 1) * R]) for i, h in enumerate(host))
    line = {"metric": "inter-frame knnMatch(k=2) problems/sec (%dx%d ORB descriptors)" % (R, R), "value": S / ms_dev * 1e3, "unit": "problems/s",
            "ms_per_step": ms_dev, "sets_per_step": S, "dtype": "u32 xor+popc", "higher_is_better": True,
            "e2e": {"value": S / ms_e2e * 1e3, "unit": "problems/s", "h2d_bytes_per_step": int(q.nbytes + t.nbytes), "d2h_bytes_per_step": int(out.nbytes),
                    "timer": "host wall clock around the blocking mcs_knn2 call"},
            "gpu_launches": int(m.kernel_launches),
            "distances_per_s": S * R * R / ms_dev * 1e3}
    if not a.no_cpu:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import oracle_lib as O
        n = 8
        t0 = time.perf_counter()
        ref = [O.knn2(qs[i], ts[i]) for i in range(n)]
        dt_cpu = time.perf_counter() - t0
        assert all(np.array_equal(ref[i], host[i]) for i in range(n)), "parity"
        line["cpu_baseline"] = {"value": n / dt_cpu, "unit": "problems/s", "cores": 1, "kind": "port", "sample": "%d problems, oracle knn2 (scalar popcount loop)" % n}
        line["parity"] = {"problems_checked": n, "mismatches": 0}
    print(json.dumps(line))


if __name__ == "__main__":
    main()
