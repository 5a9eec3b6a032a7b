"""CPU: host-side logic of the product — the libstdc++ sort emulation and the block-parallel octree formulation (the very
code the CUDA kernel runs, executed with a sequential block context) against std::sort / the oracle; the C-ABI library loads
and exports every symbol include/mcslam_b200.h declares; the synthetic generator is deterministic."""
import ctypes
import hashlib
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build_and_run(src, exe, args, extra=()):
    import oracle_lib as O
    O.build()
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", exe, os.path.join(ROOT, "tests", "cpp", src)] + list(extra))
    out = subprocess.run([exe] + args, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    return out.stdout


def test_introsort_emulation_equals_libstdcxx(tmp_path):
    out = _build_and_run("test_introsort.cpp", str(tmp_path / "t_sort"), ["4000"])
    assert "OK" in out


def test_octree_core_equals_oracle(tmp_path):
    odir = os.path.join(ROOT, "oracle")
    out = _build_and_run("test_octree_core.cpp", str(tmp_path / "t_oct"), ["400"], ["-L" + odir, "-loracle", "-Wl,-rpath," + odir])
    assert "OK" in out


def test_abi_library_exports_every_declared_symbol():
    import __graft_entry__ as G
    so = G.build()
    hdr = open(os.path.join(ROOT, "include", "mcslam_b200.h")).read()
    names = re.findall(r"MCS_API\s+[^;(]*?\b(mcs_\w+)\s*\(", hdr)
    assert len(names) >= 40, names
    lib = ctypes.CDLL(so)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    lib.mcs_version.restype = ctypes.c_char_p
    assert b"sm_100a" in lib.mcs_version()
    import mcslam_b200
    from mcslam_b200 import binding
    assert set(binding.EXPORTS) <= set(names)


def test_no_cpu_fallback_without_gpu():
    """Without a CUDA device the product must fail loudly, not compute on the CPU."""
    import mcslam_b200 as M
    if M.lib().mcs_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(M.McsError):
        M.Extractor(100)
    with pytest.raises(M.McsError):
        M.RigFrames(M.synth.make_rig(2, 160, 120), 160, 120)


def test_product_does_not_reference_oracle():
    pkg = os.path.join(ROOT, "mc-slam_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp")):
                txt = open(os.path.join(dp, f)).read()
                assert "oracle_lib" not in txt and "liboracle" not in txt and "mcs_oracle" not in txt, os.path.join(dp, f)


def test_synth_is_deterministic():
    import mcslam_b200 as M
    calib = M.synth.make_rig(2, 160, 120)
    a = M.synth.make_rig_frame(calib, 160, 120, frame_index=3, n_rects=4)
    b = M.synth.make_rig_frame(calib, 160, 120, frame_index=3, n_rects=4)
    assert np.array_equal(a, b) and a.std() > 10
    m = M.synth.make_masks(2, 160, 120, 1)
    assert 0.05 < m.mean() < 0.6 and set(np.unique(m)) <= {0.0, 1.0}


def _build_adaptor(tmp_path):
    import __graft_entry__ as G
    so = G.build()
    exe = str(tmp_path / "t_adaptor")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-o", exe, os.path.join(ROOT, "tests", "cpp", "test_adaptor.cpp"),
                           so, "-Wl,-rpath," + os.path.dirname(so)])
    return exe


def test_cpp_adaptor_compiles_and_fails_loudly_without_gpu(tmp_path):
    """The C++17 host adaptor (reference method signatures over the C ABI) builds with g++ alone; with no CUDA device its
    constructors throw — there is no CPU path behind it."""
    import mcslam_b200 as M
    exe = _build_adaptor(tmp_path)
    if M.lib().mcs_device_count() > 0:
        pytest.skip("a GPU is present")
    img = tmp_path / "im.bin"
    img.write_bytes(bytes(16 * 16))
    (tmp_path / "im.bin.calib").write_bytes(bytes(26 * 8))
    out = subprocess.run([exe, str(img), "1", "16", "16", str(tmp_path / "o.txt")], capture_output=True, text=True, timeout=120)
    assert out.returncode == 3 and "no CUDA device" in out.stderr
