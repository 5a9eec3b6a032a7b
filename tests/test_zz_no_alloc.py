"""GPU: mcs_knn2_device_static — the knn2 kernel with its scratch (the per-set offset table) in a file-scope __device__ array, so the
call allocates nothing (include/mcslam_b200.h; no counterpart in the reference) —
against the existing mcs_knn2_device on the same device buffers, twice in a row with different layouts.  Indices and distances are
integers: bit-exact, like every other knn2 test.  Named test_zz_* so that pytest collects it last."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_knn2_device_static_equals_knn2_device_twice_in_a_row():
    import mcslam_b200 as M
    from mcslam_b200 import binding as B
    rng = np.random.default_rng(77)
    q = rng.integers(0, 256, (3000, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (5000, 32), dtype=np.uint8)
    t[100:140] = t[0:40]  # ties: the lower train index must come first
    q[:40] = t[:40]
    dq, dt = B.device_malloc(0, q.nbytes), B.device_malloc(0, t.nbytes)
    d_ref, d_new = B.device_malloc(0, 3000 * 16), B.device_malloc(0, 3000 * 16)
    B.device_upload(0, dq, q)
    B.device_upload(0, dt, t)
    m = M.Matcher(max_sets=8)
    # second layout has FEWER sets than the first: stale entries of the shared table beyond n_sets must not matter
    layouts = [([1000, 0, 500, 1500], [2000, 300, 1, 2699]), ([2999, 1], [5000, 0]), ([700, 800, 1500], [1025, 2, 3973])]
    for qc, tc in layouts:
        junk = np.full((3000, 4), -7, np.int32)
        B.device_upload(0, d_ref, junk)
        B.device_upload(0, d_new, junk)
        assert m.knn2_device(dq, qc, dt, tc, d_ref) == 3000
        m.synchronize()
        assert B.knn2_device_static(0, dq, qc, dt, tc, d_new) == 3000
        ref, new = np.zeros((3000, 4), np.int32), np.zeros((3000, 4), np.int32)
        B.device_download(0, ref, d_ref)
        B.device_download(0, new, d_new)  # cudaMemcpy on the default stream: ordered after the kernels of the call above
        assert np.array_equal(new, ref), (qc, tc)
        assert (new[:, 0] >= -1).all() and (new != -7).any()
    # limits are refused with a clear error, nothing is launched
    with pytest.raises(M.McsError, match="n_sets"):
        B.knn2_device_static(0, dq, [1] * 257, dt, [1] * 257, d_new)
    with pytest.raises(M.McsError, match="65535"):
        B.knn2_device_static(0, dq, [10], dt, [65536], d_new)
    for p in (dq, dt, d_ref, d_new):
        B.device_free(0, p)
