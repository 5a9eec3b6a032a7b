// Host-side C++ adaptor smoke: mirrors how FrontEnd drives the reference classes (FrontEnd.cpp:991-1028).
//   argv: <raw u8 images file [C][h][w]> C w h <out file>   -> writes counts, keypoints, descriptors, rows, landmarks as text
// Without a GPU the constructors throw (no CPU path): exit code 3.
#include <cstdio>
#include <cstdlib>
#include <fstream>

#include "../../mc-slam_b200/host/mcslam_adaptor.hpp"

int main(int argc, char** argv) {
    if (argc < 6) return 2;
    const int C = atoi(argv[2]), w = atoi(argv[3]), h = atoi(argv[4]);
    std::vector<uint8_t> img((size_t)C * w * h);
    std::ifstream(argv[1], std::ios::binary).read((char*)img.data(), img.size());
    std::vector<mcs_calib> calib(C);
    std::ifstream(std::string(argv[1]) + ".calib", std::ios::binary).read((char*)calib.data(), C * sizeof(mcs_calib));
    try {
        mcs_b200::ORBextractor ex(400, 1.2f, 8, 20, 7, 0, w, h);
        std::vector<mcs_keypoint> k;
        std::vector<uint8_t> d;
        std::vector<int> lap = {0, 0};
        const int mono = ex(img.data(), w, h, w, nullptr, k, d, lap);
        FILE* f = fopen(argv[5], "w");
        fprintf(f, "mono %d n %zu levels %d sf %.9g\n", mono, k.size(), ex.GetLevels(), ex.GetScaleFactors()[1]);
        for (size_t i = 0; i < k.size(); i++) fprintf(f, "kp %.9g %.9g %d %.9g %d\n", k[i].x, k[i].y, k[i].octave, k[i].response, d[i * 32]);
        mcs_b200::MultiCameraFrame fr(calib, w, h, 400, 1.2f, 8, 20, 7);
        fr.setData(img.data());
        fr.extractFeaturesParallel();
        std::vector<mcs_b200::IntraMatch> rows, lms;
        fr.computeIntraMatches(rows, false);
        fr.triangulateIntraMatches((const float*)nullptr, lms);
        fprintf(f, "rows %zu landmarks %zu\n", rows.size(), lms.size());
        for (auto& r : rows) fprintf(f, "row %d %d %d\n", r.matchIndexN[0], r.matchIndexN[1], r.matchIndexN[C - 1]);
        for (auto& m : lms) fprintf(f, "lm %d %.12g %.12g %.12g\n", m.n_rays, m.point3D[0], m.point3D[1], m.point3D[2]);
        fclose(f);
        uint8_t empty = 0;
        if (ex(nullptr, 0, 0, 0, nullptr, k, d, lap) != -1) return 4;  // `if(_image.empty()) return -1;`
        (void)empty;
    } catch (const std::exception& e) {
        fprintf(stderr, "adaptor: %s\n", e.what());
        return 3;
    }
    return 0;
}
