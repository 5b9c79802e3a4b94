// Drives include/viorb_fextractor.hpp the way Frame::ExtractORB does (reference
// src/datastructures/frame.cpp:416-426) and dumps the results for tests/test_cpp_shim.py.
// usage: shim_driver <in.raw> <w> <h> <nfeatures> <lap0> <lap1> <out.bin>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include "viorb_fextractor.hpp"

using vi_slam::geometry::FExtractor;

int main(int argc, char** argv) {
    if (argc != 8) return 2;
    const int w = std::atoi(argv[2]), h = std::atoi(argv[3]), nf = std::atoi(argv[4]);
    std::vector<int> lap = {std::atoi(argv[5]), std::atoi(argv[6])};
    std::vector<uint8_t> buf((size_t)w * h);
    FILE* f = std::fopen(argv[1], "rb");
    if (!f || std::fread(buf.data(), 1, buf.size(), f) != buf.size()) return 3;
    std::fclose(f);
    // two instances entered concurrently from two threads, as frame.cpp:107-111 does for left / right
    FExtractor a(nf, 1.2f, 8, 20, 7, w, h), b(nf, 1.2f, 8, 20, 7, w, h);
    viorb_cv::Mat im(h, w, buf.data(), (size_t)w), mask;
    std::vector<viorb_cv::KeyPoint> ka, kb;
    viorb_cv::Mat da, db;
    int ma = -2, mb = -2;
    std::thread ta([&] { ma = a.compute(im, mask, ka, da, lap); });
    std::thread tb([&] { mb = b.compute(im, mask, kb, db, lap); });
    ta.join(); tb.join();
    if (ma != mb || ka.size() != kb.size() || da.rows != db.rows) return 4;
    if (!ka.empty() && std::memcmp(ka.data(), kb.data(), ka.size() * sizeof(ka[0]))) return 5;
    viorb_cv::Mat empty;
    if (a.compute(empty, mask, kb, db, lap) != -1) return 6;               // fextractor.cpp:1037
    if (a.GetLevels() != 8 || a.GetScaleFactors().size() != 8 || a.mvImagePyramid.size() != 8) return 7;
    FILE* o = std::fopen(argv[7], "wb");
    const int n = (int)ka.size();
    std::fwrite(&ma, 4, 1, o); std::fwrite(&n, 4, 1, o);
    if (n) { std::fwrite(ka.data(), sizeof(ka[0]), n, o); std::fwrite(da.data, 32, n, o); }
    for (int l = 0; l < 8; ++l) {
        const viorb_cv::Mat& m = a.mvImagePyramid[l];
        std::fwrite(&m.cols, 4, 1, o); std::fwrite(&m.rows, 4, 1, o);
        std::fwrite(m.data, 1, (size_t)m.cols * m.rows, o);
    }
    std::fclose(o);
    return 0;
}
