// Compiles include/of_b200_shim.hpp against a minimal cv::Mat-shaped type and mirrors the
// reference's pyramidal harness (ParallelBasicPyr.cpp:717-836): reads two raw f64 images, calls
// getMaxLayer + coarseToFineEstimation + getLucasKanadeOpticalFlow by their reference names,
// writes u, v as raw f64.  Usage: shim_check W H img1.f64 img2.f64 out_prefix [nLevels]
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "of_b200_shim.hpp"

struct MiniMat {                      // the subset of cv::Mat the shim touches
    int rows = 0, cols = 0;
    unsigned char* data = nullptr;
    size_t step = 0;
    std::vector<double> store;
    int type() const { return OF_B200_CV_64FC1; }
    bool empty() const { return rows == 0 || cols == 0; }
    void create(int r, int c, int)
    {
        rows = r; cols = c; step = (size_t)c * sizeof(double);
        store.assign((size_t)r * c, 0.0);
        data = reinterpret_cast<unsigned char*>(store.data());
    }
};

static bool load(MiniMat& m, int w, int h, const char* path)
{
    m.create(h, w, OF_B200_CV_64FC1);
    FILE* f = fopen(path, "rb");
    if (!f) return false;
    const size_t n = fread(m.store.data(), sizeof(double), (size_t)w * h, f);
    fclose(f);
    return n == (size_t)w * h;
}

static bool save(const MiniMat& m, const std::string& path)
{
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) return false;
    fwrite(m.store.data(), sizeof(double), m.store.size(), f);
    fclose(f);
    return true;
}

int main(int argc, char** argv)
{
    if (argc < 6) { fprintf(stderr, "usage: %s W H img1 img2 out_prefix [nLevels]\n", argv[0]); return 2; }
    const int w = atoi(argv[1]), h = atoi(argv[2]);
    MiniMat img1, img2, u, v, u1, v1;
    if (!load(img1, w, h, argv[3]) || !load(img2, w, h, argv[4])) { fprintf(stderr, "cannot read inputs\n"); return 2; }
    int maxLayer = getMaxLayer(img1);                                  // ParallelBasicPyr.cpp:726
    if (argc > 6) maxLayer = atoi(argv[6]);
    printf("%d, %d, Max layer = %d\n", h, w, maxLayer);
    try {
        coarseToFineEstimation(img1, img2, u, v, maxLayer);            // :753
        getLucasKanadeOpticalFlow(img1, img2, u1, v1);                 // serialOrig.cpp:559
    } catch (const std::exception& e) {
        fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
    const std::string p = argv[5];
    return save(u, p + "_u.f64") && save(v, p + "_v.f64") && save(u1, p + "_u1.f64") && save(v1, p + "_v1.f64") ? 0 : 2;
}
