// of_b200_shim.hpp -- header-only C++ shim that re-exposes the reference's own entry points
// over the C ABI of of_b200.h, so a reference main() keeps compiling with one #include and one
// link flag changed.  Reference signatures (ParallelVideoPyr.cpp):
//
//   int  getMaxLayer(Mat &img)                                                   :545
//   void getLucasKanadeOpticalFlow(Mat &img1, Mat &img2, Mat &u, Mat &v)         :327
//   void coarseToFineEstimation(Mat &img1, Mat &img2, Mat &u, Mat &v, int nLevels) :578
//
// The shim is a template over the matrix type so that it does not need OpenCV headers: any type
// with the cv::Mat members used below works (rows, cols, data, step, type(), create(), empty()).
// With OpenCV present, MatT = cv::Mat and OF_B200_CV_64FC1 = CV_64FC1 (= 6).
//
// Behaviour kept from the reference: u and v are (re)allocated by the callee (u = tmpu rebinding,
// :639-640; callers pass 0x0 Mats, :822-823); a size not divisible by 2^(nLevels-1) throws (the
// reference throws cv::Exception at tmpu += upu, :635).  Behaviour NOT kept by default: the
// reference overwrites img1/img2 with their warped versions (quirk Q1, :696,701); construct the
// context with write_back_warped = 1 for that.
#pragma once
#include <stdexcept>
#include <string>

#include "of_b200.h"

#ifndef OF_B200_CV_64FC1
#define OF_B200_CV_64FC1 6
#endif

namespace of_b200 {

class Context {
public:
    explicit Context(int device = 0, const of_params* params = nullptr)
    {
        if (of_ctx_create(device, params, &ctx_) != OF_OK)
            throw std::runtime_error(std::string("of_ctx_create: ") + of_last_error_string(nullptr));
    }
    ~Context() { of_ctx_destroy(ctx_); }
    Context(const Context&) = delete;
    Context& operator=(const Context&) = delete;
    of_ctx* get() const { return ctx_; }

    // One process-wide context on device 0 with the reference's constants, created on first use.
    static Context& instance()
    {
        static Context c(0, nullptr);
        return c;
    }

private:
    of_ctx* ctx_ = nullptr;
};

inline void check(of_ctx* ctx, of_status st, const char* what)
{
    if (st == OF_OK) return;
    const std::string msg = std::string(what) + ": " + of_last_error_string(ctx);
    if (st == OF_E_BADARG || st == OF_E_NOT_DIVISIBLE) throw std::invalid_argument(msg);
    throw std::runtime_error(msg);
}

template <class MatT>
void require_f64(const MatT& m, const char* name)
{
    if (m.empty() || m.type() != OF_B200_CV_64FC1)
        throw std::invalid_argument(std::string(name) + " must be a non-empty CV_64FC1 matrix");
}

}  // namespace of_b200

// int getMaxLayer(Mat &img)   ParallelVideoPyr.cpp:545-564
template <class MatT>
inline int getMaxLayer(MatT& img)
{
    return of_get_max_layer(img.cols, img.rows);
}

// void getLucasKanadeOpticalFlow(Mat&, Mat&, Mat&, Mat&)   ParallelVideoPyr.cpp:327-542
template <class MatT>
inline void getLucasKanadeOpticalFlow(MatT& img1, MatT& img2, MatT& u, MatT& v)
{
    of_b200::require_f64(img1, "img1");
    of_b200::require_f64(img2, "img2");
    if (img1.rows != img2.rows || img1.cols != img2.cols || (size_t)img1.step != (size_t)img2.step)
        throw std::invalid_argument("img1 and img2 must have the same size and step");
    u.create(img1.rows, img1.cols, OF_B200_CV_64FC1);
    v.create(img1.rows, img1.cols, OF_B200_CV_64FC1);
    of_ctx* ctx = of_b200::Context::instance().get();
    of_b200::check(ctx,
                   of_lk_single_f64(ctx, reinterpret_cast<const double*>(img1.data),
                                    reinterpret_cast<const double*>(img2.data), img1.cols, img1.rows,
                                    (size_t)img1.step, reinterpret_cast<double*>(u.data),
                                    reinterpret_cast<double*>(v.data), (size_t)u.step, nullptr, 0),
                   "getLucasKanadeOpticalFlow");
}

// void coarseToFineEstimation(Mat&, Mat&, Mat&, Mat&, int)   ParallelVideoPyr.cpp:578-725
template <class MatT>
inline void coarseToFineEstimation(MatT& img1, MatT& img2, MatT& u, MatT& v, int nLevels)
{
    of_b200::require_f64(img1, "img1");
    of_b200::require_f64(img2, "img2");
    if (img1.rows != img2.rows || img1.cols != img2.cols || (size_t)img1.step != (size_t)img2.step)
        throw std::invalid_argument("img1 and img2 must have the same size and step");
    u.create(img1.rows, img1.cols, OF_B200_CV_64FC1);
    v.create(img1.rows, img1.cols, OF_B200_CV_64FC1);
    of_ctx* ctx = of_b200::Context::instance().get();
    of_b200::check(ctx,
                   of_ctf_f64(ctx, reinterpret_cast<double*>(img1.data), reinterpret_cast<double*>(img2.data),
                              img1.cols, img1.rows, (size_t)img1.step, nLevels,
                              reinterpret_cast<double*>(u.data), reinterpret_cast<double*>(v.data),
                              (size_t)u.step, nullptr, 0),
                   "coarseToFineEstimation");
}
