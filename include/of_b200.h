/* of_b200.h -- C ABI of the B200-native dense pyramidal Lucas-Kanade path.
 *
 * This is the drop-in boundary.  The reference
 * (mileslindheimer/optical-flow) has no plugin/FFI interface: its boundary is the set
 * of C++ free functions every main() calls.  Each entry point below names the
 * reference function it replaces (file:line into the reference repository; line
 * numbers are ParallelVideoPyr.cpp's, serial twin ctfLKof.cpp:121-215).  A header-only
 * C++ shim with the reference's own function names lives in of_b200_shim.hpp.
 *
 * Layout conventions (== the reference's cv::Mat CV_64FC1 convention):
 *   frames and flow planes are single-channel, row-major double; element (y,x) is at
 *   (char*)base + y*step_bytes + x*8.  step_bytes may exceed width*8 (ROI views) and
 *   must be a multiple of 8.  u = x-flow, v = y-flow, two separate planes.
 *   mask (optional, may be NULL): uint8 plane, 1 where the level-0 2x2 system passed
 *   the validity rule (|det| > det_eps; det != 0 with the reference's det_eps = 0).
 *
 * All functions are extern "C", POD-only, and never throw.  A context is bound to one
 * CUDA device and is not thread-safe; distinct contexts may be used concurrently from
 * distinct threads/processes (the multi-GPU model: one process or thread per GPU; the
 * of_mgpu_* calls below do exactly that inside one process).
 * Streams: all calls on one context share its device scratch (pyramids, per-level flows).
 * Calls are ordered against each other whatever stream they are given: a call on a stream
 * other than the previous call's waits (on the device) for that call's kernels first.
 * There is NO CPU fallback: if no CUDA device is usable, of_ctx_create fails with
 * OF_E_CUDA.
 */
#ifndef OF_B200_H_
#define OF_B200_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OF_B200_ABI_VERSION 1

#if defined(__GNUC__)
#define OF_API __attribute__((visibility("default")))
#else
#define OF_API
#endif

typedef struct of_ctx of_ctx;
typedef int of_status;

enum {
    OF_OK = 0,
    OF_E_BADARG = 1,        /* null pointer, width/height < 1, n_levels < 1, bad step  */
    OF_E_NOT_DIVISIBLE = 2, /* width or height % 2^(n_levels-1) != 0: the reference      */
                            /* throws a cv::Exception at tmpu += upu (:635), quirk Q2    */
    OF_E_CUDA = 3,          /* CUDA runtime/driver error; see of_last_error_string       */
    OF_E_NOMEM = 4          /* device or host allocation failed                          */
};

typedef struct of_params {
    int win_radius;        /* window radius r, window = (2r+1)^2.  Reference: 1 (get_Sum9,   */
                           /* :114-160).  BASELINE throughput config: 7 (15x15).  1..15.     */
    double det_eps;        /* validity: |det| > det_eps.  Reference: 0.0 (cv::divide's        */
                           /* zero-divisor -> 0 rule, :514,518).                              */
    int write_back_warped; /* 1: host-pointer of_ctf_f64 overwrites img1/img2 with their      */
                           /* warped level-0 versions like the reference does (quirk Q1,      */
                           /* warped.copyTo(pyr[i-1]) :696,701).  Default 0: inputs const.    */
                           /* The _dev, batch and stream calls take const inputs and return   */
                           /* OF_E_BADARG when the flag is 1 and n_levels > 1.                */
    double mag_thresh;     /* gesture statistics: flow magnitude threshold, reference 20 (:859)*/
    int count_thresh;      /* gesture present iff count > count_thresh, reference 500 (:886)  */
} of_params;

/* Per-pair gesture statistics (ParallelVideoPyr.cpp:845-888, headless replacement of
 * circle()+imshow()): sums of x (column) and y (row) indices and the count of pixels
 * whose flow magnitude is not < mag_thresh. */
typedef struct of_gesture_stats {
    double sum_x;
    double sum_y;
    double count;     /* integer-valued */
    double detected;  /* 1.0 iff count > count_thresh */
} of_gesture_stats;

/* Fills *p with the reference's constants (r=1, det_eps=0, no write-back, 20, 500). */
OF_API void of_params_default(of_params *p);

/* Replaces: int getMaxLayer(Mat &img)   ParallelVideoPyr.cpp:545-564 (ctfLKof.cpp:196-215).
 * min over width/height of 1 + (number of times 2 divides n).  Returns 0 for empty sizes. */
OF_API int of_get_max_layer(int width, int height);

/* Context: owns a CUDA stream and all device scratch (pyramids, per-level flows, staging
 * buffers); scratch is sized on first use per (width, height, n_levels, batch) and reused.
 * params may be NULL (reference defaults). */
OF_API of_status of_ctx_create(int device, const of_params *params, of_ctx **out);
OF_API void of_ctx_destroy(of_ctx *ctx);
OF_API of_status of_ctx_set_params(of_ctx *ctx, const of_params *params);
OF_API const char *of_last_error_string(const of_ctx *ctx); /* ctx may be NULL: last create error */
OF_API int of_abi_version(void);

/* ---- host-pointer entry points (synchronous; H2D + compute + D2H) ----------------------- */

/* Replaces: void getLucasKanadeOpticalFlow(Mat &img1, Mat &img2, Mat &u, Mat &v)
 *           ParallelVideoPyr.cpp:327-542 (ctfLKof.cpp:121-144).  Single level, no pyramid. */
OF_API of_status of_lk_single_f64(of_ctx *ctx, const double *img1, const double *img2,
                           int width, int height, size_t step_bytes,
                           double *u, double *v, size_t uv_step_bytes,
                           unsigned char *mask, size_t mask_step_bytes);

/* Replaces: void coarseToFineEstimation(Mat &img1, Mat &img2, Mat &u, Mat &v, int nLevels)
 *           ParallelVideoPyr.cpp:578-725 (ctfLKof.cpp:157-194), including
 *           getGaussianPyramid :566-575.  img1/img2 are written only when
 *           params.write_back_warped = 1 (hence non-const). */
OF_API of_status of_ctf_f64(of_ctx *ctx, double *img1, double *img2,
                     int width, int height, size_t step_bytes, int n_levels,
                     double *u, double *v, size_t uv_step_bytes,
                     unsigned char *mask, size_t mask_step_bytes);

/* Frame stream, host pointers: n_frames frames, frame k at (char*)frames + k*frame_stride_bytes;
 * computes the n_frames-1 consecutive pairs (k, k+1) (the live loop's pairing,
 * ParallelVideoPyr.cpp:831-843, with inputs read-only) and writes flow k at
 * (char*)u + k*uv_pair_stride_bytes.  u/v may be NULL when only stats are wanted; stats may
 * be NULL.  Uploads are chunked and overlapped with compute on the context's streams. */
OF_API of_status of_ctf_f64_stream(of_ctx *ctx, const double *frames, int n_frames,
                            size_t frame_stride_bytes, int width, int height,
                            size_t step_bytes, int n_levels,
                            double *u, double *v, size_t uv_pair_stride_bytes,
                            size_t uv_step_bytes, of_gesture_stats *stats);

/* The same stream call for 8-bit frames (what imread(..., 0) / VideoCapture deliver): each frame is
 * uploaded as bytes and converted on the device, pixel = rn(rn(v * alpha) + beta).
 * Replaces: imread(...,0) frames + img.convertTo(img, CV_64FC1, 1.0/255)  ParallelBasicPyr.cpp:720-730
 * (alpha = 1.0/255, beta = 0) followed by coarseToFineEstimation :753 per consecutive pair.
 * step_bytes / frame_stride_bytes are in bytes of the 8-bit frames; 1/8 of the host-to-device
 * traffic of of_ctf_f64_stream. */
OF_API of_status of_ctf_u8_stream(of_ctx *ctx, const unsigned char *frames, int n_frames,
                           size_t frame_stride_bytes, int width, int height,
                           size_t step_bytes, double alpha, double beta, int n_levels,
                           double *u, double *v, size_t uv_pair_stride_bytes,
                           size_t uv_step_bytes, of_gesture_stats *stats);

/* ---- device-pointer entry points (stream-ordered, asynchronous) -------------------------- */
/* All pointers are device pointers on the context's device.  stream is a cudaStream_t passed
 * as void* (NULL = the context's own stream).  These are what the benchmark's device-resident
 * number times. */

OF_API of_status of_lk_single_f64_dev(of_ctx *ctx, const double *img1, const double *img2,
                               int width, int height, size_t step_bytes,
                               double *u, double *v, size_t uv_step_bytes,
                               unsigned char *mask, size_t mask_step_bytes, void *stream);

OF_API of_status of_ctf_f64_dev(of_ctx *ctx, const double *img1, const double *img2,
                         int width, int height, size_t step_bytes, int n_levels,
                         double *u, double *v, size_t uv_step_bytes,
                         unsigned char *mask, size_t mask_step_bytes, void *stream);

/* Batch of independent pairs: pair k reads img1 + k*img_pair_stride_bytes and
 * img2 + k*img_pair_stride_bytes, writes u/v + k*uv_pair_stride_bytes.  A video stream is the
 * special case img2 = img1 + one frame, img_pair_stride = one frame.  stats (device pointer to
 * n_pairs of_gesture_stats, may be NULL) receives the gesture statistics of every pair;
 * u and v may both be NULL when only stats are wanted. */
OF_API of_status of_ctf_f64_batch_dev(of_ctx *ctx, int n_pairs,
                               const double *img1, const double *img2,
                               size_t img_pair_stride_bytes,
                               int width, int height, size_t step_bytes, int n_levels,
                               double *u, double *v, size_t uv_pair_stride_bytes,
                               size_t uv_step_bytes,
                               unsigned char *mask, size_t mask_pair_stride_bytes,
                               size_t mask_step_bytes,
                               of_gesture_stats *stats, void *stream);

/* Gesture statistics of existing flow planes (ParallelVideoPyr.cpp:845-873). */
OF_API of_status of_gesture_stats_dev(of_ctx *ctx, int n_pairs, const double *u, const double *v,
                               size_t uv_pair_stride_bytes, int width, int height,
                               size_t uv_step_bytes, of_gesture_stats *stats, void *stream);

/* Building blocks, exposed for parity triage (each is one reference/OpenCV call site). */
/* cv::pyrDown at getGaussianPyramid :571.  dst is ((w+1)/2) x ((h+1)/2). */
OF_API of_status of_pyr_down_f64_dev(of_ctx *ctx, const double *src, int width, int height,
                              size_t step_bytes, double *dst, size_t dst_step_bytes, void *stream);

/* ---- live-loop pre-processing (the reference's main(), ParallelVideoPyr.cpp:791-813) ---- */

/* Replaces: cvtColor(current_frame, current_frame, CV_BGR2GRAY)   ParallelVideoPyr.cpp:791.
 * 8-bit interleaved BGR -> 8-bit gray, fixed point.  coeff_bits selects OpenCV's table:
 * 14 = OpenCV 2.4 (the reference's library: 1868, 9617, 4899), 15 = OpenCV >= 3 (3735, 19235, 9798).
 * Device pointers. */
OF_API of_status of_bgr2gray_u8_dev(of_ctx *ctx, const unsigned char *bgr, int width, int height,
                             size_t bgr_step_bytes, unsigned char *gray, size_t gray_step_bytes,
                             int coeff_bits, void *stream);

/* Replaces the block between the gray frame and the LK input of one live-loop iteration:
 *   Mat diff = current_frame - LEARNING_RATE * prevFrame            :800  (8-bit, saturating)
 *   threshold(diff, diff, DIFF_THRESH, 255, THRESH_TOZERO)           :803
 *   Sobel(diff, .., CV_64FC1, 1, 0) + Sobel(diff, .., CV_64FC1, 0, 1) :808-811
 *   dilate(diff, diff, Mat(), Point(-1,-1), 2); erode(.., 2)          :812-813
 * cur_gray / prev_gray: 8-bit gray frames of the same size and step; out: double plane (the img2
 * of this iteration's coarseToFineEstimation call, :832).  Reference constants: learning_rate 0.3,
 * diff_thresh 10, morph_iterations 2 (1..4 supported).  Device pointers; one fused kernel. */
OF_API of_status of_live_diff_f64_dev(of_ctx *ctx, const unsigned char *cur_gray,
                               const unsigned char *prev_gray, int width, int height,
                               size_t gray_step_bytes, double learning_rate, int diff_thresh,
                               int morph_iterations, double *out, size_t out_step_bytes,
                               void *stream);

/* Host-pointer form of the two calls above (synchronous): cur / prev are 8-bit frames with
 * `channels` = 1 (gray) or 3 (BGR, converted with coeff_bits first), out a host double plane. */
OF_API of_status of_live_diff_f64(of_ctx *ctx, const unsigned char *cur, const unsigned char *prev,
                           int channels, int width, int height, size_t frame_step_bytes,
                           int coeff_bits, double learning_rate, int diff_thresh,
                           int morph_iterations, double *out, size_t out_step_bytes);

/* Replaces: resize(frame, current_frame, Size(1080, 1080), 0, 0, INTER_CUBIC)   ParallelVideoPyr.cpp:782.
 * 8-bit, `channels` = 1 or 3 interleaved; cv::resize's fixed-point bicubic (A = -0.75, 11-bit
 * coefficients, BORDER_REPLICATE taps) as cv2 4.13 computes it without IPP.  Equal sizes copy.  Device
 * pointers; the coefficient tables are cached in the context per geometry. */
OF_API of_status of_resize_cubic_u8_dev(of_ctx *ctx, const unsigned char *src, int src_width, int src_height,
                                 size_t src_step_bytes, int channels, unsigned char *dst, int dst_width,
                                 int dst_height, size_t dst_step_bytes, void *stream);

/* Replaces: GaussianBlur(current_frame, current_frame, Size(9, 9), 1.5, 1.5)   ParallelVideoPyr.cpp:790
 * (+ optionally cvtColor(CV_BGR2GRAY) :791 fused: gray_coeff_bits = 14 / 15 with channels = 3 writes one
 * gray channel; 0 writes the blurred image).  8-bit fixed point: integer kernel / 256 in both passes
 * (for 9 / 1.5: 2 9 28 55 68 55 28 9 2), BORDER_REFLECT_101, (sum + 2^15) >> 16.  ksize odd, <= 31.
 * src and dst must not overlap.  Device pointers. */
OF_API of_status of_gaussian_blur_u8_dev(of_ctx *ctx, const unsigned char *src, int width, int height,
                                  size_t src_step_bytes, int channels, int ksize, double sigma,
                                  int gray_coeff_bits, unsigned char *dst, size_t dst_step_bytes, void *stream);

/* One iteration of the reference's live loop, headless (main(), ParallelVideoPyr.cpp:769-905, without
 * VideoCapture / imshow / circle): the captured 8-bit BGR frame (host pointer) goes through
 *   resize :782 -> GaussianBlur :790 -> cvtColor :791 -> cur - 0.3 prev :800 -> threshold :803 ->
 *   Sobel x + y :808-811 -> dilate x2 / erode x2 :812-813 -> coarseToFineEstimation(prevDiff, diff) :832
 *   -> gesture statistics :845-888
 * entirely on the device; the context keeps prevFrame / prevDiff between calls like the loop does.
 * *stage = 0: first frame stored (:793-797), 1: first difference image stored (:816-820), 2: *stats is
 * valid (and u, v when given: host planes of work_width x work_height doubles, may be NULL).
 * With of_params.write_back_warped = 1 prevDiff becomes the WARPED difference image, as in the reference
 * (quirk Q1: coarseToFineEstimation overwrites its inputs, then prevDiff = diff :843); default: unwarped. */
typedef struct of_live_params {
    int work_width, work_height; /* 1080, 1080 (:782)                                               */
    int blur_ksize;              /* 9 (:790)                                                        */
    double blur_sigma;           /* 1.5                                                             */
    int coeff_bits;              /* cvtColor table: 14 = OpenCV 2.4 (the reference's), 15 = OpenCV >= 3 */
    double learning_rate;        /* 0.3  (LEARNING_RATE :729)                                       */
    int diff_thresh;             /* 10   (DIFF_THRESH :728)                                         */
    int morph_iterations;        /* 2    (:812-813)                                                 */
    int n_levels;                /* 0 = getMaxLayer(work size) (:831), else the given pyramid depth */
} of_live_params;
OF_API void of_live_params_default(of_live_params *p);
OF_API of_status of_live_step_u8(of_ctx *ctx, const of_live_params *lp, const unsigned char *frame_bgr,
                          int width, int height, size_t step_bytes, of_gesture_stats *stats,
                          double *u, double *v, size_t uv_step_bytes, int *stage);
OF_API of_status of_live_reset(of_ctx *ctx);   /* forget prevFrame / prevDiff: the next frame is a first frame */

/* Blocks until everything queued on the context's own stream has finished. */
OF_API of_status of_ctx_synchronize(of_ctx *ctx);

/* Per-launch device timing (bench roofline evidence).  When enabled, every kernel of a batch
 * call is bracketed by CUDA events on the launching stream (up to 8192 launches since the last
 * of_ctx_set_timing call).  of_ctx_get_timing returns, for those launches in order, kinds[i] (0 = pyrDown producing level levels[i], 1 = fused level
 * kernel of level levels[i]) and the device duration ms[i]; it waits for those events. */
OF_API of_status of_ctx_set_timing(of_ctx *ctx, int enable);
OF_API of_status of_ctx_get_timing(of_ctx *ctx, int max_entries, int *kinds, int *levels,
                                   float *ms, int *n);

/* Parity triage: copies one plane of pyramid level `level` >= 1 as left in the
 * context by the most recent pyramid call: the Gaussian pyramid image of frame 1 / frame 2 of
 * pair `pair` (getGaussianPyramid :566-575; unwarped -- warped levels never exist in memory) or
 * the accumulated flow u_l / v_l of that level (tmpu / tmpv after :634-637).  dst may be a host
 * or a device pointer; dst = NULL only reports the level's size.  Waits for that call. */
enum { OF_LEVEL_FRAME1 = 0, OF_LEVEL_FRAME2 = 1, OF_LEVEL_FLOW_U = 2, OF_LEVEL_FLOW_V = 3 };
OF_API of_status of_ctx_get_level_f64(of_ctx *ctx, int what, int level, int pair, double *dst,
                                      size_t dst_step_bytes, int *width, int *height);

/* ---- single-process multi-GPU (frame pairs shard across the GPUs of one box) ---------
 * The reference is one C++ process; this is the C/C++ way to put its frame stream on several GPUs:
 * one context per device, one host thread per device inside each call, the n_frames - 1 consecutive
 * pairs split into contiguous blocks (of_mgpu_partition), flow fields and statistics written straight
 * into the caller's arrays.  A frame is never split and the devices exchange nothing (no collective).
 * devices = NULL / n_devices = 0: every visible device.  Arguments as of_ctf_f64_stream /
 * of_ctf_u8_stream (host pointers; pinned memory makes the copies asynchronous). */
typedef struct of_mgpu of_mgpu;
OF_API of_status of_mgpu_create(const int *devices, int n_devices, const of_params *params, of_mgpu **out);
OF_API void of_mgpu_destroy(of_mgpu *m);
OF_API int of_mgpu_device_count(const of_mgpu *m);
OF_API of_ctx *of_mgpu_context(of_mgpu *m, int index);   /* the context of the index-th device (set_params, timing) */
OF_API const char *of_mgpu_last_error_string(const of_mgpu *m); /* m may be NULL: last create error */
/* Pairs [*first, *last) go to the index-th of n_devices devices. */
OF_API of_status of_mgpu_partition(int n_pairs, int n_devices, int index, int *first, int *last);
OF_API of_status of_mgpu_ctf_f64_stream(of_mgpu *m, const double *frames, int n_frames,
                                 size_t frame_stride_bytes, int width, int height,
                                 size_t step_bytes, int n_levels,
                                 double *u, double *v, size_t uv_pair_stride_bytes,
                                 size_t uv_step_bytes, of_gesture_stats *stats);
OF_API of_status of_mgpu_ctf_u8_stream(of_mgpu *m, const unsigned char *frames, int n_frames,
                                size_t frame_stride_bytes, int width, int height,
                                size_t step_bytes, double alpha, double beta, int n_levels,
                                double *u, double *v, size_t uv_pair_stride_bytes,
                                size_t uv_step_bytes, of_gesture_stats *stats);

/* ---- allocation-free variant (no counterpart in the reference)
 * ------------------------------------------------------------------------------------------
 * of_ctf_f64_dev for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it): the same kernels and results for ONE pair, with the device scratch
 * (pyramid levels >= 1 of both frames, flows of the levels >= 1; about 10.7 bytes per level-0 pixel) in a
 * static device array of OF_STATIC_SCRATCH_BYTES inside the library instead of the context's arena.  The
 * call allocates nothing (no cudaMalloc / cudaFree / allocating library routine); every GPU that loads
 * the library holds the array until the process exits.  Sizes that need more scratch are rejected with
 * OF_E_BADARG and a message naming both numbers (32 MiB covers up to 1920x1080 at any depth).  The array
 * is shared by every call on a device and keeps the last call's contents; nothing in it has to start at
 * zero.  Calls that overlap on one GPU race on it: serialise them (one stream, or events between them).
 * All pointers are device pointers; stream as in of_ctf_f64_dev; no mask, no statistics; rejects
 * write_back_warped like the other _dev calls. */
#define OF_STATIC_SCRATCH_BYTES (32u << 20)
OF_API of_status of_ctf_f64_dev_static(of_ctx *ctx, const double *img1, const double *img2,
                                int width, int height, size_t step_bytes, int n_levels,
                                double *u, double *v, size_t uv_step_bytes, void *stream);

/* Number of kernel launches issued through this context since creation (bench evidence). */
OF_API unsigned long long of_ctx_launch_count(const of_ctx *ctx);

/* Small batches (up to 4 Mpixel per call: the harness pair, one 1080p pair -- the reference's own call
 * pattern, one coarseToFineEstimation per frame, ParallelVideoPyr.cpp:831-832) are launch-bound.  The
 * second time a call repeats exactly (same geometry, parameters and pointers) its launch sequence is
 * captured into a CUDA graph and replayed from then on; kernels launched through a replay count in
 * of_ctx_launch_count.  This returns the number of replays (environment OF_B200_GRAPH=0 disables them). */
OF_API unsigned long long of_ctx_graph_replay_count(const of_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* OF_B200_H_ */
