"""Builds oracle/_ref/libmcsref.so: the REFERENCE's own front-end code compiled here as the parity checker.

TEST INFRASTRUCTURE ONLY (tests/, __graft_entry__.smoke(), bench.py's cpu_baseline / --impl reference).

What is compiled, from where it lies under /root/reference (nothing is copied into the repository; the extracted ranges go to the
git-ignored oracle/_ref/gen/):
  * MCSlam/src/ORBextractor.cpp — the whole file, unmodified;
  * the method bodies listed in RANGES, cut by line range from MultiCameraFrame.cpp, FrontEnd.cpp, MultiCameraFrame.h, utils.h,
    tools.cpp and included verbatim by oracle/ref_shim/ref_bodies.cpp inside class shells (oracle/ref_shim/ref_shells.hpp);
  * MCDataUtils/include/MCDataUtils/CamArrayConfig.h and MCSlam/include/MCSlam/ORBextractor.h as they are.
OpenCV / Eigen / DBoW2 / glog are absent from this image: oracle/ref_shim/mini_cv.hpp stands in for the OpenCV API surface and
forwards the pixel primitives to liboracle.so's cv2-pinned restatements.  Each range is checked against the first and last line
it is expected to hold, so a different reference checkout fails loudly instead of compiling something else.

On the GPU box /root/reference does not exist: the prebuilt oracle/_ref/libmcsref.so travels with the snapshot.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("MCS_REFERENCE_ROOT", "/root/reference")
OUT_DIR = os.path.join(HERE, "_ref")
GEN = os.path.join(OUT_DIR, "gen")
SO = os.path.join(OUT_DIR, "libmcsref.so")
SHIM = os.path.join(HERE, "ref_shim")

# (output, file, first line, last line, text the first line must contain, text the last line must equal after strip)
RANGES = [
    ("IntraMatch.inc", "MCSlam/include/MCSlam/MultiCameraFrame.h", 41, 57, "// todo: generalize for n cameras", "};"),
    ("utils_argsorte.inc", "MCSlam/include/MCSlam/utils.h", 20, 30, "template<typename T>", "}"),
    ("tools_build_Rt.inc", "common_utils/src/tools.cpp", 136, 148, "Mat build_Rt(Mat R, Mat t) {", "}"),
    ("mcf_extractFeatures.inc", "MCSlam/src/MultiCameraFrame.cpp", 203, 262, "void MultiCameraFrame::extractFeaturesParallel(){", "}"),
    ("mcf_UndistortKeyPoints.inc", "MCSlam/src/MultiCameraFrame.cpp", 300, 347, "void MultiCameraFrame::UndistortKeyPoints(int cam) {", "}"),
    ("mcf_triangulateIntraMatches.inc", "MCSlam/src/MultiCameraFrame.cpp", 411, 441,
     "void MultiCameraFrame::triangulateIntraMatches(const Mat_<double> &x, vector<int> view_inds, Eigen::Vector3d& pt){", "}"),
    ("mcf_triangulate.inc", "MCSlam/src/MultiCameraFrame.cpp", 480, 507, "void MultiCameraFrame::", "}"),
    ("mcf_computeRepresentativeDesc.inc", "MCSlam/src/MultiCameraFrame.cpp", 530, 567,
     "void MultiCameraFrame::computeRepresentativeDesc( vector<Mat> descs, cv::Mat& desc_out ){", "}"),
    ("mcf_BruteForceMatch.inc", "MCSlam/src/MultiCameraFrame.cpp", 1024, 1086, "void MultiCameraFrame::BruteForceMatch(int img1_ind, int img2_ind,", "}"),
    ("mcf_computeIntraMatches.inc", "MCSlam/src/MultiCameraFrame.cpp", 1100, 1288,
     "void MultiCameraFrame:: computeIntraMatches( vector<IntraMatch>& matches, bool old){", "}"),
    ("frontend_obtainLfFeatures.inc", "MCSlam/src/FrontEnd.cpp", 213, 593, "void FrontEnd::obtainLfFeatures(std::vector<IntraMatch>& matches_map,", "}"),
]

CXXFLAGS = ["-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-march=x86-64-v3", "-pthread", "-fvisibility=default", "-w"]


def reference_present():
    return os.path.isfile(os.path.join(REF, "MCSlam", "src", "ORBextractor.cpp"))


def extract():
    os.makedirs(GEN, exist_ok=True)
    for out, rel, a, b, head, tail in RANGES:
        with open(os.path.join(REF, rel), encoding="utf-8", errors="replace") as f:
            lines = f.readlines()
        chunk = lines[a - 1:b]
        if head not in chunk[0] or chunk[-1].strip() != tail:
            raise RuntimeError("%s:%d-%d does not hold the expected text (first %r, last %r)" % (rel, a, b, chunk[0], chunk[-1]))
        with open(os.path.join(GEN, out), "w", encoding="utf-8") as f:
            f.write("// extracted verbatim from /root/reference/%s:%d-%d by oracle/ref_build.py - build output, do not commit\n" % (rel, a, b))
            f.write('#line %d "%s"\n' % (a, os.path.join(REF, rel)))
            f.writelines(chunk)


def sources():
    return [os.path.join(REF, "MCSlam", "src", "ORBextractor.cpp")] + [os.path.join(SHIM, f) for f in ("ref_bodies.cpp", "ref_abi.cpp", "mini_cv.cpp")]


def needs_build():
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = sources() + [os.path.join(SHIM, f) for f in ("mini_cv.hpp", "ref_shells.hpp")] + [os.path.join(HERE, "liboracle.so"), __file__]
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force=False, verbose=False):
    """Returns the path of libmcsref.so, building it when /root/reference is present and it is stale; None when it cannot exist."""
    if not reference_present():
        return SO if os.path.exists(SO) else None
    if not force and not needs_build():
        return SO
    subprocess.check_call(["make", "-C", HERE, "-s", "liboracle.so"])
    extract()
    cxx = os.environ.get("CXX", "g++")
    cmd = [cxx] + CXXFLAGS + ["-shared", "-o", SO, "-I", SHIM, "-I", OUT_DIR, "-I", os.path.join(REF, "MCSlam", "include"),
                              "-I", os.path.join(REF, "MCDataUtils", "include")] + sources() + [
        "-L", HERE, "-loracle", "-Wl,-rpath,$ORIGIN/.."]
    if verbose:
        print(" ".join(cmd))
    try:
        subprocess.check_call(cmd)
    finally:
        shutil.rmtree(GEN, ignore_errors=True)  # the extracted reference text is needed for the compile only
    return SO


if __name__ == "__main__":
    so = build(force="-B" in sys.argv, verbose=True)
    print(so if so else "reference sources not present and no prebuilt oracle/_ref/libmcsref.so")
