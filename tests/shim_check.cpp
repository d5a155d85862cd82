// Compiles include/of_b200_shim.hpp against a minimal cv::Mat-shaped type and mirrors the
// reference's pyramidal harness (ParallelBasicPyr.cpp:717-836): reads two raw f64 images, calls
// getMaxLayer + coarseToFineEstimation + getLucasKanadeOpticalFlow by their reference names,
// writes u, v as raw f64.  Usage: shim_check W H img1.f64 img2.f64 out_prefix [nLevels]
//
// Second mode, the C++ caller's way onto several GPUs (of_mgpu_*, one process):
//   shim_check --mgpu W H n_frames frames.f64 nLevels radius
// runs the frame stream on every visible GPU through of_mgpu_ctf_f64_stream and on GPU 0 alone through
// of_ctf_f64_stream, and requires identical flow fields and statistics.  Prints "mgpu ok <n_devices>".
#include <cstdio>
#include <cstdlib>
#include <cstring>
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

static int mgpu_mode(int argc, char** argv)
{
    if (argc < 8) { fprintf(stderr, "usage: %s --mgpu W H n_frames frames.f64 nLevels radius\n", argv[0]); return 2; }
    const int w = atoi(argv[2]), h = atoi(argv[3]), n = atoi(argv[4]), L = atoi(argv[6]);
    const size_t px = (size_t)w * h;
    std::vector<double> fr(px * n);
    FILE* f = fopen(argv[5], "rb");
    if (!f || fread(fr.data(), 8, fr.size(), f) != fr.size()) { fprintf(stderr, "cannot read frames\n"); return 2; }
    fclose(f);
    of_params p;
    of_params_default(&p);
    p.win_radius = atoi(argv[7]);
    p.mag_thresh = 0.5;
    p.count_thresh = 100;
    std::vector<double> u1(px * (n - 1)), v1(u1.size()), um(u1.size()), vm(u1.size());
    std::vector<of_gesture_stats> s1(n - 1), sm(n - 1);
    of_ctx* ctx = nullptr;
    if (of_ctx_create(0, &p, &ctx) != OF_OK) { fprintf(stderr, "error: %s\n", of_last_error_string(nullptr)); return 1; }
    of_status st = of_ctf_f64_stream(ctx, fr.data(), n, px * 8, w, h, (size_t)w * 8, L, u1.data(), v1.data(), px * 8,
                                     (size_t)w * 8, s1.data());
    if (st != OF_OK) { fprintf(stderr, "error: %s\n", of_last_error_string(ctx)); return 1; }
    of_ctx_destroy(ctx);
    of_mgpu* m = nullptr;
    if (of_mgpu_create(nullptr, 0, &p, &m) != OF_OK) { fprintf(stderr, "error: %s\n", of_mgpu_last_error_string(nullptr)); return 1; }
    st = of_mgpu_ctf_f64_stream(m, fr.data(), n, px * 8, w, h, (size_t)w * 8, L, um.data(), vm.data(), px * 8, (size_t)w * 8,
                                sm.data());
    if (st != OF_OK) { fprintf(stderr, "error: %s\n", of_mgpu_last_error_string(m)); return 1; }
    const int nd = of_mgpu_device_count(m);
    of_mgpu_destroy(m);
    if (memcmp(u1.data(), um.data(), u1.size() * 8) || memcmp(v1.data(), vm.data(), v1.size() * 8) ||
        memcmp(s1.data(), sm.data(), s1.size() * sizeof(of_gesture_stats))) {
        fprintf(stderr, "multi-GPU result differs from the single-GPU result\n");
        return 1;
    }
    printf("mgpu ok %d\n", nd);
    return 0;
}

int main(int argc, char** argv)
{
    if (argc > 1 && !strcmp(argv[1], "--mgpu")) return mgpu_mode(argc, argv);
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
