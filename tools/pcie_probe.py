import sys, time, numpy as np
sys.path.insert(0, ".")
import mcslam_b200 as M
from mcslam_b200 import binding as B
n = 512 << 20
hp = B.pinned_empty((n,), np.uint8); h = hp.arr
h[:] = 1
d = B.device_malloc(0, n)
for _ in range(2): B.device_upload(0, d, h)
t0 = time.perf_counter()
for _ in range(5): B.device_upload(0, d, h)
t = (time.perf_counter() - t0) / 5
print("H2D pinned GB/s", n / t / 1e9)
t0 = time.perf_counter()
for _ in range(5): B.device_download(0, h, d)
t = (time.perf_counter() - t0) / 5
print("D2H pinned GB/s", n / t / 1e9)
