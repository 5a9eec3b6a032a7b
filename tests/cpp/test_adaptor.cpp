// Host-side C++ adaptor smoke: mirrors how FrontEnd drives the reference classes (FrontEnd.cpp:991-1028).
//   argv: <raw u8 images file [C][h][w]> C w h <out file>   -> writes counts, keypoints, descriptors, rows, landmarks as text
//   argv: inter <dir> <out file>                              -> inter-frame matchers + pairwise triangulation (see run_inter)
// Without a GPU the constructors throw (no CPU path): exit code 3.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <string>

#include "../../mc-slam_b200/host/mcslam_adaptor.hpp"

template <typename T>
static std::vector<T> slurp(const std::string& path) {
    std::ifstream f(path, std::ios::binary | std::ios::ate);
    std::vector<T> v((size_t)f.tellg() / sizeof(T));
    f.seekg(0);
    f.read((char*)v.data(), v.size() * sizeof(T));
    return v;
}

// argv: inter <dir> <out>   -- FrontEnd::findMatchesMono / findInterMatches (both overloads) and the pairwise
// triangulateIntraMatches(kps_1, kps_2) through the adaptor, on arrays the python test wrote into <dir>.
static int run_inter(const std::string& dir, const char* out_path) {
    using namespace mcs_b200;
    int n1, n2, w, isz1, isz2, C, npair;
    FILE* meta = fopen((dir + "/meta.txt").c_str(), "r");
    if (!meta || fscanf(meta, "%d %d %d %d %d %d %d", &n1, &n2, &w, &isz1, &isz2, &C, &npair) != 7) return 2;
    fclose(meta);
    auto d1 = slurp<uint8_t>(dir + "/d1.bin"), d2 = slurp<uint8_t>(dir + "/d2.bin");
    auto xy1 = slurp<float>(dir + "/xy1.bin"), xy2 = slurp<float>(dir + "/xy2.bin");
    auto m1 = slurp<float>(dir + "/m1.bin"), m2 = slurp<float>(dir + "/m2.bin");
    auto mono1 = slurp<uint8_t>(dir + "/mono1.bin"), mono2 = slurp<uint8_t>(dir + "/mono2.bin");
    auto p1 = slurp<double>(dir + "/p1.bin"), p2 = slurp<double>(dir + "/p2.bin");
    auto lids = slurp<int>(dir + "/lids1.bin");
    FILE* f = fopen(out_path, "w");
    InterFrameMatcher im(0, 4096);
    std::vector<mcs_keypoint> k1(n1), k2(n2);
    for (int i = 0; i < n1; i++) { k1[i] = mcs_keypoint{}; k1[i].x = xy1[2 * i]; k1[i].y = xy1[2 * i + 1]; }
    for (int i = 0; i < n2; i++) { k2[i] = mcs_keypoint{}; k2[i].x = xy2[2 * i]; k2[i].y = xy2[2 * i + 1]; }
    std::vector<DMatch> mm;
    im.findMatchesMono(d1.data(), k1, m1.data(), d2.data(), k2, m2.data(), w, mm);
    for (auto& m : mm) fprintf(f, "mono %d %d %d\n", m.queryIdx, m.trainIdx, (int)m.distance);
    std::vector<IntraMatch> a(n1), b(n2);
    for (int i = 0; i < n1; i++) {
        memcpy(a[i].matchDesc.data(), &d1[(size_t)i * 32], 32);
        a[i].mono = mono1[i] != 0;
        for (int k = 0; k < 3; k++) a[i].point3D[k] = p1[3 * i + k];
    }
    for (int i = 0; i < n2; i++) {
        memcpy(b[i].matchDesc.data(), &d2[(size_t)i * 32], 32);
        b[i].mono = mono2[i] != 0;
        for (int k = 0; k < 3; k++) b[i].point3D[k] = p2[3 * i + k];
    }
    std::vector<int> none(n2, -1);
    InterFrameView prev{&a, &lids, isz1}, cur{&b, &none, isz2};
    std::vector<DMatch> lf, mo, single;
    im.findInterMatches(prev, cur, lf, mo);
    im.findInterMatches(prev, cur, single);
    for (auto& m : lf) fprintf(f, "lf %d %d %d\n", m.queryIdx, m.trainIdx, (int)m.distance);
    for (auto& m : mo) fprintf(f, "mo %d %d %d\n", m.queryIdx, m.trainIdx, (int)m.distance);
    for (auto& m : single) fprintf(f, "single %d %d %d\n", m.queryIdx, m.trainIdx, (int)m.distance);
    // empty train set: no rows, no crash (the reference would index m[0] of an empty vector)
    std::vector<IntraMatch> nobody;
    InterFrameView empty{&nobody, &none, 0};
    im.findInterMatches(prev, empty, single);
    fprintf(f, "empty %zu\n", single.size());

    // pairwise triangulation wrapper: pairs.bin = for every camera pair (i < j) npair rows of (x1 y1 x2 y2) f32
    std::vector<mcs_calib> calib(C);
    std::ifstream(dir + "/calib.bin", std::ios::binary).read((char*)calib.data(), C * sizeof(mcs_calib));
    auto pairs = slurp<float>(dir + "/pairs.bin");
    MultiCameraFrame fr(calib, 320, 240, 300, 1.2f, 8, 20, 7);
    std::vector<std::vector<std::vector<std::array<float, 2>>>> kp1(C - 1), kp2(C - 1);
    size_t o = 0;
    for (int i = 0; i < C - 1; i++)
        for (int j = i + 1; j < C; j++) {
            std::vector<std::array<float, 2>> u(npair), v(npair);
            for (int k = 0; k < npair; k++, o += 4) { u[k] = {pairs[o], pairs[o + 1]}; v[k] = {pairs[o + 2], pairs[o + 3]}; }
            kp1[i].push_back(u);
            kp2[i].push_back(v);
        }
    {   // BoW-guided matcher: bow.bin = node id per feature (A then B), descriptors bowA.bin / bowB.bin
        auto bA = slurp<uint8_t>(dir + "/bowA.bin"), bB = slurp<uint8_t>(dir + "/bowB.bin");
        auto nodes = slurp<int>(dir + "/bow_nodes.bin");
        const int nA = (int)bA.size() / 32, nB = (int)bB.size() / 32;
        FeatureVector f1, f2;
        for (int i = 0; i < nA; i++) f1[(unsigned)nodes[i]].push_back((unsigned)i);
        for (int i = 0; i < nB; i++) f2[(unsigned)nodes[nA + i]].push_back((unsigned)i);
        BowMatcher bm(0, 4096);
        std::vector<unsigned int> i1, i2;
        std::set<unsigned int> words;
        bm.BowMatching(f1, f2, bA.data(), nA, bB.data(), nB, i1, i2, &words);
        for (size_t k = 0; k < i1.size(); k++) fprintf(f, "bow %u %u\n", i1[k], i2[k]);
        for (unsigned wd : words) fprintf(f, "word %u\n", wd);
        int book = 0;
        const unsigned first = f1.begin()->first;
        if (f2.count(first)) {
            bm.getMatches_distRatio(bA.data(), nA, f1[first], bB.data(), nB, f2[first], i1, i2, book);
            for (size_t k = 0; k < i1.size(); k++) fprintf(f, "node0 %u %u\n", i1[k], i2[k]);
            fprintf(f, "book %d\n", book);
        }
    }
    {   // vocabulary: voc.txt in <dir>, transform of the bowA descriptors at levelsup 2
        ORBVocabulary voc;
        if (voc.loadFromTextFile(dir + "/voc.txt")) {
            auto bA = slurp<uint8_t>(dir + "/bowA.bin");
            FeatureVector fv;
            voc.transform(bA.data(), (int)bA.size() / 32, fv, 2);
            for (auto& kv : fv) {
                fprintf(f, "fv %u", kv.first);
                for (unsigned i : kv.second) fprintf(f, " %u", i);
                fprintf(f, "\n");
            }
        }
    }
    fr.triangulateIntraMatches(kp1, kp2);
    for (auto& p : fr.pts3D_) fprintf(f, "pt %.9g %.9g %.9g\n", p[0], p[1], p[2]);
    fclose(f);
    return 0;
}

int main(int argc, char** argv) {
    if (argc >= 4 && std::string(argv[1]) == "inter") {
        try {
            return run_inter(argv[2], argv[3]);
        } catch (const std::exception& e) {
            fprintf(stderr, "adaptor: %s\n", e.what());
            return 3;
        }
    }
    if (argc >= 6 && std::string(argv[1]) == "calib") {  // calib <yaml> <radtan 0|1> <camchain 0|1> <out>: the Kalibr camchain reader
        try {
            mcs_b200::CamArrayConfig cfg = mcs_b200::read_kalibr_data(argv[2], atoi(argv[3]) != 0, atoi(argv[4]) != 0);
            FILE* f = fopen(argv[5], "w");
            fprintf(f, "size %d %d %d\n", cfg.num_cams(), cfg.width, cfg.height);
            for (int c = 0; c < cfg.num_cams(); c++) {
                fprintf(f, "topic %s\n", cfg.cam_topics[c].c_str());
                const double* v = (const double*)&cfg.calib[c];
                fprintf(f, "cam");
                for (int k = 0; k < 26; k++) fprintf(f, " %.17g", v[k]);
                fprintf(f, "\nraw");
                v = (const double*)&cfg.calib_kalibr[c];
                for (int k = 0; k < 26; k++) fprintf(f, " %.17g", v[k]);
                fprintf(f, "\n");
            }
            fclose(f);
            return 0;
        } catch (const std::exception& e) {
            fprintf(stderr, "adaptor: %s\n", e.what());
            return 3;
        }
    }
    if (argc >= 7 && std::string(argv[1]) == "overlap") {  // overlap <calib bin> C w h <out>: FrontEnd::compute_overlap (host, no GPU)
        const int C = atoi(argv[3]), w = atoi(argv[4]), h = atoi(argv[5]);
        std::vector<mcs_calib> calib(C);
        std::ifstream(argv[2], std::ios::binary).read((char*)calib.data(), C * sizeof(mcs_calib));
        const mcs_b200::OverlapInfo o = mcs_b200::compute_overlap(calib, w, h);
        FILE* f = fopen(argv[6], "wb");
        for (int c = 0; c < C; c++) fwrite(o.overlap_masks[c].data(), 1, o.overlap_masks[c].size(), f);
        fwrite(o.non_overlap_fraction.data(), sizeof(float), C, f);
        fclose(f);
        return 0;
    }
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
        mcs_b200::MultiCameraFrame::LfFeatures lf;
        fr.obtainLfFeatures(nullptr, lf, 700);  // small budget so that the response cut is exercised
        fprintf(f, "lfsizes %d %d %zu\n", lf.intramatch_size, lf.mono_size, lf.lIds.size());
        for (int i = lf.intramatch_size; i < (int)lf.intraMatches.size(); i++) {
            const auto& m = lf.intraMatches[i];
            int c = 0;
            while (m.matchIndexN[c] < 0) c++;
            fprintf(f, "lfmono %d %d %d %d\n", c, m.matchIndexN[c], (int)m.mono, m.matchDesc[0]);
        }
        {   // BoW-guided computeIntraMatches(matches, words_): vocabulary in <images>.voc.txt, FeatureVectors at levelsup argv[6]
            mcs_b200::ORBVocabulary voc;
            if (argc >= 7 && voc.loadFromTextFile(std::string(argv[1]) + ".voc.txt")) {
                fr.BoW_feats.assign(C, mcs_b200::FeatureVector());
                for (int c = 0; c < C; c++)
                    voc.transform(fr.image_descriptors[c].data(), (int)fr.image_descriptors[c].size() / 32, fr.BoW_feats[c], atoi(argv[6]));
                std::vector<mcs_b200::IntraMatch> bm;
                std::vector<unsigned int> words;
                fr.computeIntraMatches(bm, words);
                for (auto& r : bm) {
                    fprintf(f, "bowrow %d", r.n_rays);
                    for (int c = 0; c < C; c++) fprintf(f, " %d", r.matchIndexN[c]);
                    fprintf(f, "\n");
                }
                for (unsigned w : words) fprintf(f, "bowword %u\n", w);
                std::vector<mcs_b200::IntraMatch> chained;
                fr.computeIntraMatches_chained(chained, false);
                for (auto& r : chained) {
                    fprintf(f, "chainrow");
                    for (int c = 0; c < C; c++) fprintf(f, " %d", r.matchIndexN[c]);
                    fprintf(f, "\n");
                }
                std::vector<std::vector<std::vector<mcs_b200::DMatch>>> dm;
                std::vector<std::vector<std::vector<std::array<float, 2>>>> k1, k2;
                fr.computeIntraMatches(dm, k1, k2);
                for (size_t i = 0; i < dm.size(); i++)
                    for (size_t j = 0; j < dm[i].size(); j++)
                        for (size_t k = 0; k < dm[i][j].size(); k++)
                            fprintf(f, "dm %zu %zu %d %d %.9g %.9g %.9g %.9g\n", i, i + 1 + j, dm[i][j][k].queryIdx, dm[i][j][k].trainIdx,
                                    k1[i][j][k][0], k1[i][j][k][1], k2[i][j][k][0], k2[i][j][k][1]);
            }
        }
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
