// Test driver for include/LynxHardwareAccelerator.h (the C++ drop-in of ORB_SLAM3::LynxHardwareAccelerator): feeds
// `nframes` side-by-side frames through StoreInputBuffer + ExtractORB, then asks BFMatchORB for the cached matches of
// every frame id AFTER all extractions (only the last MATCHING_CACHE_SIZE ids are still cached), and dumps everything
// for tests/test_gpu_wire.py.  Exit code 3 = the constructor threw (no CUDA device): there is no fallback.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "LynxHardwareAccelerator.h"

static void dump(FILE* f, int count, int mono, const std::vector<cv::KeyPoint>& k, const cv::Mat& d) {
    const int n = (int)k.size();
    fwrite(&n, 4, 1, f);
    fwrite(&count, 4, 1, f);
    fwrite(&mono, 4, 1, f);
    if (n) fwrite(k.data(), sizeof(cv::KeyPoint), n, f);
    if (n) fwrite(d.data, 32, n, f);
}

int main(int argc, char** argv) {
    if (argc < 11) {
        fprintf(stderr, "usage: %s frames.raw w h nframes nfeatures lapL0 lapL1 lapR0 lapR1 out.bin\n", argv[0]);
        return 2;
    }
    const int w = atoi(argv[2]), h = atoi(argv[3]), nframes = atoi(argv[4]), nf = atoi(argv[5]);
    const int lap[4] = {atoi(argv[6]), atoi(argv[7]), atoi(argv[8]), atoi(argv[9])};
    std::vector<unsigned char> frames((size_t)2 * w * h * nframes);
    FILE* fi = fopen(argv[1], "rb");
    if (!fi || fread(frames.data(), 1, frames.size(), fi) != frames.size()) return 2;
    fclose(fi);
    try {
        ORB_SLAM3::LynxHardwareAccelerator lha(w, h, nf, 1.2f, 8, 20, 7);
        FILE* fo = fopen(argv[10], "wb");
        std::vector<int> ids, stereoRows;
        for (int f = 0; f < nframes; f++) {
            lha.StoreInputBuffer(&frames[(size_t)f * 2 * w * h]);
            int nl = 0, nr = 0, ml = 0, mr = 0;
            std::vector<cv::KeyPoint> kl, kr;
            cv::Mat dl, dr;
            const int id = lha.ExtractORB(nl, nr, kl, kr, dl, dr, lap[0], lap[1], lap[2], lap[3], ml, mr);
            fwrite(&id, 4, 1, fo);
            dump(fo, nl, ml, kl, dl);
            dump(fo, nr, mr, kr, dr);
            ids.push_back(id);
            stereoRows.push_back(nl - ml);
        }
        for (int f = 0; f < nframes; f++) {    // Frame.cc:1164: BFMatchORB(mnIdMatchingData, stereoDescRight, stereoDescLeft, ...)
            cv::Mat stereoDescRight, stereoDescLeft(stereoRows[f], 32, CV_8U);
            std::vector<uint16_t> indices, dist1, dist2;
            lha.BFMatchORB(ids[f], stereoDescRight, stereoDescLeft, indices, dist1, dist2);
            const int n = (int)indices.size();
            fwrite(&n, 4, 1, fo);
            if (n) {
                fwrite(indices.data(), 2, n, fo);
                fwrite(dist1.data(), 2, n, fo);
                fwrite(dist2.data(), 2, n, fo);
            }
        }
        fclose(fo);
    } catch (const std::exception& e) {
        fprintf(stderr, "%s\n", e.what());
        return 3;
    }
    return 0;
}
