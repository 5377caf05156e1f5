// Test driver for include/ORBextractor.h (the C++ drop-in class): reads a raw 8-bit image, calls the classic
// ORBextractor::operator() (and the stereo overload on a side-by-side frame built from two copies of the input) and
// dumps the results for the Python tests.  Exit code 3 = the constructor threw (no CUDA device): there is no fallback.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "ORBextractor.h"

static void dump(FILE* f, int mono, const std::vector<cv::KeyPoint>& k, const cv::Mat& d) {
    const int n = (int)k.size();
    fwrite(&n, 4, 1, f);
    fwrite(&mono, 4, 1, f);
    if (n) fwrite(k.data(), sizeof(cv::KeyPoint), n, f);
    if (n) fwrite(d.data, 32, n, f);
}

int main(int argc, char** argv) {
    if (argc < 9) {
        fprintf(stderr, "usage: %s in.raw w h nfeatures nlevels lap0 lap1 out.bin\n", argv[0]);
        return 2;
    }
    const int w = atoi(argv[2]), h = atoi(argv[3]), nf = atoi(argv[4]), nl = atoi(argv[5]);
    std::vector<int> lap = {atoi(argv[6]), atoi(argv[7])};
    std::vector<unsigned char> img((size_t)w * h);
    FILE* fi = fopen(argv[1], "rb");
    if (!fi || fread(img.data(), 1, img.size(), fi) != img.size()) return 2;
    fclose(fi);
    try {
        ORB_SLAM3::ORBextractor ex(nf, 1.2f, nl, 20, 7);
        cv::Mat image(h, w, CV_8UC1, img.data()), mask, desc;
        std::vector<cv::KeyPoint> kps;
        const int mono = ex(image, mask, kps, desc, lap);
        FILE* fo = fopen(argv[8], "wb");
        dump(fo, mono, kps, desc);
        // getters + pyramid
        const std::vector<float> sf = ex.GetScaleFactors();
        fwrite(sf.data(), 4, sf.size(), fo);
        for (int l = 0; l < ex.GetLevels(); l++) {
            const cv::Mat& p = ex.mvImagePyramid[l];
            const int dims[2] = {p.cols, p.rows};
            fwrite(dims, 4, 2, fo);
            fwrite(p.data, 1, (size_t)p.rows * p.cols, fo);
        }
        // stereo overload: left = the image, right = the image again
        std::vector<unsigned char> sbs((size_t)2 * w * h);
        for (int y = 0; y < h; y++) {
            memcpy(&sbs[(size_t)y * 2 * w], &img[(size_t)y * w], w);
            memcpy(&sbs[(size_t)y * 2 * w + w], &img[(size_t)y * w], w);
        }
        std::vector<cv::KeyPoint> kl, kr;
        cv::Mat dl, dr;
        int ml = 0, mr = 0;
        std::vector<int> lapR = {0, 0};
        const int id = ex(sbs.data(), w, h, 2 * w, kl, dl, lap, kr, dr, lapR, ml, mr);
        fwrite(&id, 4, 1, fo);
        dump(fo, ml, kl, dl);
        dump(fo, mr, kr, dr);
        fclose(fo);
    } catch (const std::exception& e) {
        fprintf(stderr, "%s\n", e.what());
        return 3;
    }
    return 0;
}
