// C-ABI layer (include/of_b200.h) over the sm_100a kernels: context, device scratch,
// level scheduling of coarseToFineEstimation (ParallelVideoPyr.cpp:578-725) and the
// host-pointer / device-pointer / batch / stream entry points.
//
// There is no CPU path in this file: every entry point either launches the CUDA kernels
// or returns an error code.
#include <nvtx3/nvToolsExt.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "of_kernels.h"

namespace ofb {
cudaError_t launch_lk_level_tile(const LevelArgs& a, int n_pairs, cudaStream_t s,
                                 unsigned long long* launches);
cudaError_t launch_lk_level_phase(const LevelArgs& a, int n_pairs, cudaStream_t s,
                                  unsigned long long* launches, bool* handled);
}  // namespace ofb

namespace {

std::string g_create_error;

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct Arena {
    char* base = nullptr;
    size_t cap = 0;
};

}  // namespace

// A captured launch sequence, keyed on everything that ends up in a kernel parameter.
struct GraphKey {
    int n_pairs, w, h, L, radius, count_thresh;
    double det_eps, mag_thresh;
    const void *img1, *img2, *u, *v, *mask, *stats;
    size_t img_pair_stride, step, uv_pair_stride, uv_step, mask_pair_stride, mask_step;
    const void* scratch;
    bool operator==(const GraphKey& o) const
    {
        return n_pairs == o.n_pairs && w == o.w && h == o.h && L == o.L && radius == o.radius &&
               count_thresh == o.count_thresh && det_eps == o.det_eps && mag_thresh == o.mag_thresh && img1 == o.img1 &&
               img2 == o.img2 && u == o.u && v == o.v && mask == o.mask && stats == o.stats &&
               img_pair_stride == o.img_pair_stride && step == o.step && uv_pair_stride == o.uv_pair_stride &&
               uv_step == o.uv_step && mask_pair_stride == o.mask_pair_stride && mask_step == o.mask_step &&
               scratch == o.scratch;
    }
};
struct GraphEntry {
    GraphKey key{};
    cudaGraphExec_t exec = nullptr;
    bool seen = false;
    unsigned long long launches = 0, last_use = 0;
};

struct of_ctx {
    int device = 0;
    of_params params{};
    cudaStream_t stream = nullptr;     // compute stream owned by the context
    cudaStream_t copy_stream = nullptr;  // H2D uploads of the stream entry point
    cudaStream_t down_stream = nullptr;  // D2H downloads of the stream entry point
    cudaEvent_t ev_up[2] = {nullptr, nullptr};
    cudaEvent_t ev_done[2] = {nullptr, nullptr};
    cudaEvent_t ev_down[2] = {nullptr, nullptr};
    Arena scratch;                     // pyramids + per-level flows
    Arena staging;                     // device copies of host-pointer inputs/outputs
    Arena live;                        // live loop: frame, resized frame, gray frames, difference images
    // resize tables (device), cached per geometry
    struct ResizeTab { int sw = 0, sh = 0, dw = 0, dh = 0; int *xofs = nullptr, *yofs = nullptr; short* xco = nullptr; float* ybeta = nullptr; } rz;
    // live-loop state (of_live_step_u8): prevFrame / prevDiff of main() :765-766
    struct LiveState { int w = 0, h = 0, frames = 0, cur_gray = 0, cur_diff = 0; } ls;
    of_gesture_stats* h_stats = nullptr;  // pinned host staging of the stream call's statistics
    size_t h_stats_cap = 0;               // (a copy into pageable memory would block the host thread)
    std::string err;
    unsigned long long launches = 0;
    std::vector<GraphEntry> graphs;    // at most 8, least recently used replaced
    unsigned long long graph_clock = 0, graph_replays = 0;
    int graph_mode = 1;                // 0 never, 1 small batches, 2 always (OF_B200_GRAPH, test hook)
    int force_variant = 0;             // 0 auto, 1 tile, 4 phase (test hook via OF_B200_VARIANT)
    // optional per-launch device timing of the most recent batch call (bench roofline evidence)
    bool timing = false;
    std::vector<cudaEvent_t> tev;      // 2 events per timed launch
    std::vector<int> tkind, tlevel;
    size_t tcount = 0;
    // geometry of the most recent batch call (of_ctx_get_level_f64, parity triage)
    struct LastLevel { int w, h; size_t pitch, img_stride, pyr_off, flow_u_off, flow_v_off; };
    std::vector<LastLevel> last;
    int last_pairs = 0;
    bool last_shared = false;
    // The scratch arena is shared by every call on this context: ev_scratch marks the end of the most
    // recent batch on the stream it ran on (last_stream), so that a call on a different stream waits
    // for it instead of overwriting pyramids / flows that are still being read.
    cudaStream_t last_stream = nullptr;
    cudaEvent_t ev_scratch = nullptr;
    bool scratch_busy = false;
};

namespace {

of_status fail(of_ctx* ctx, of_status code, const char* fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf; else g_create_error = buf;
    return code;
}

#define OF_CUDA(ctx, call)                                                                  \
    do {                                                                                    \
        cudaError_t e_ = (call);                                                            \
        if (e_ != cudaSuccess)                                                              \
            return fail(ctx, e_ == cudaErrorMemoryAllocation ? OF_E_NOMEM : OF_E_CUDA,       \
                        "%s failed: %s", #call, cudaGetErrorString(e_));                    \
    } while (0)

void graph_clear(of_ctx* ctx)
{
    for (GraphEntry& e : ctx->graphs)
        if (e.exec) cudaGraphExecDestroy(e.exec);
    ctx->graphs.clear();
}

bool graph_eligible(const of_ctx* ctx, int n_pairs, int w, int h)
{
    if (ctx->timing || ctx->graph_mode == 0) return false;       // per-launch timing needs the launches themselves
    return ctx->graph_mode == 2 || (long long)n_pairs * w * h <= (1ll << 22);
}

GraphEntry* graph_lookup(of_ctx* ctx, const GraphKey& key)
{
    ctx->graph_clock += 1;
    for (GraphEntry& e : ctx->graphs)
        if (e.key == key) {
            e.last_use = ctx->graph_clock;
            return &e;
        }
    if (ctx->graphs.size() >= 8) {
        size_t victim = 0;
        for (size_t i = 1; i < ctx->graphs.size(); ++i)
            if (ctx->graphs[i].last_use < ctx->graphs[victim].last_use) victim = i;
        if (ctx->graphs[victim].exec) cudaGraphExecDestroy(ctx->graphs[victim].exec);
        ctx->graphs.erase(ctx->graphs.begin() + victim);
    }
    GraphEntry e;
    e.key = key;
    e.last_use = ctx->graph_clock;
    ctx->graphs.push_back(e);
    return &ctx->graphs.back();
}

of_status arena_reserve(of_ctx* ctx, Arena& a, size_t bytes)
{
    if (bytes <= a.cap) return OF_OK;
    if (&a == &ctx->scratch) graph_clear(ctx);                   // captured launches point into the old arena
    if (a.base) {
        OF_CUDA(ctx, cudaDeviceSynchronize());
        OF_CUDA(ctx, cudaFree(a.base));
        a.base = nullptr;
        a.cap = 0;
    }
    bytes = align_up(bytes + bytes / 8, (size_t)1 << 20);
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(ctx, OF_E_NOMEM, "cudaMalloc(%zu bytes) failed: %s", bytes, cudaGetErrorString(e));
    }
    a.base = (char*)p;
    a.cap = bytes;
    return OF_OK;
}

struct LevelGeom {
    int w, h;
    size_t pitch;       // doubles
    size_t img_stride;  // doubles
};

// Scratch layout for one batch.
struct Plan {
    int L = 0, n_pairs = 0, n_frames = 0;  // n_frames: images per pyramid level
    bool shared = false;                   // stream mode: frame k+1 of pair k is frame k of pair k+1
    std::vector<LevelGeom> g;              // per level
    std::vector<size_t> pyr_off;           // byte offset of level l's images (l >= 1)
    std::vector<size_t> flow_u_off, flow_v_off;
    size_t total = 0;
};

// l0_flow: statistics-only calls (u = v = NULL) get scratch level-0 flow planes, so that they run on the
// same throughput kernels as calls that return the flow (those park the upsampled flow in u, v).
Plan make_plan(int w, int h, int L, int n_pairs, bool shared, bool l0_flow)
{
    Plan p;
    p.L = L;
    p.n_pairs = n_pairs;
    p.shared = shared;
    p.n_frames = shared ? n_pairs + 1 : 2 * n_pairs;
    p.g.resize(L);
    p.pyr_off.assign(L, 0);
    p.flow_u_off.assign(L, 0);
    p.flow_v_off.assign(L, 0);
    size_t off = 0;
    for (int l = 0; l < L; ++l) {
        LevelGeom& g = p.g[l];
        g.w = w >> l;
        g.h = h >> l;
        g.pitch = align_up((size_t)g.w, 16);             // 128-byte rows
        g.img_stride = align_up(g.pitch * (size_t)g.h, 32);
        if (l >= 1) {
            p.pyr_off[l] = off;
            off += g.img_stride * p.n_frames * sizeof(double);
        }
        if (l >= 1 || l0_flow) {
            p.flow_u_off[l] = off;
            off += g.img_stride * n_pairs * sizeof(double);
            p.flow_v_off[l] = off;
            off += g.img_stride * n_pairs * sizeof(double);
        }
    }
    p.total = off;
    return p;
}

of_status check_common(of_ctx* ctx, const void* a, const void* b, int w, int h, size_t step,
                       int L)
{
    if (!ctx) return OF_E_BADARG;
    if (!a || !b) return fail(ctx, OF_E_BADARG, "null frame pointer");
    if (w < 1 || h < 1) return fail(ctx, OF_E_BADARG, "width/height must be >= 1 (got %dx%d)", w, h);
    if (w > 32767 || h > 32767)
        return fail(ctx, OF_E_BADARG, "width/height above 32767 are not representable in remap's int16 coordinates");
    if (step % sizeof(double) || step < (size_t)w * sizeof(double))
        return fail(ctx, OF_E_BADARG, "step_bytes %zu invalid for width %d", step, w);
    if (L < 1 || L > 24) return fail(ctx, OF_E_BADARG, "n_levels must be in 1..24 (got %d)", L);
    const int f = 1 << (L - 1);
    if (w % f || h % f)
        return fail(ctx, OF_E_NOT_DIVISIBLE, "%dx%d is not divisible by 2^(n_levels-1) = %d", w, h, f);
    return OF_OK;
}

// One validation for of_ctx_create and of_ctx_set_params.
const char* validate_params(const of_params& p)
{
    if (p.win_radius < 1 || p.win_radius > 15) return "win_radius must be in 1..15";
    if (!(p.det_eps >= 0.0) || std::isinf(p.det_eps)) return "det_eps must be finite and >= 0";
    if (p.write_back_warped != 0 && p.write_back_warped != 1) return "write_back_warped must be 0 or 1";
    if (p.mag_thresh != p.mag_thresh) return "mag_thresh is NaN";
    if (p.count_thresh < 0) return "count_thresh must be >= 0";
    return nullptr;
}

// Smallest double s with !(sqrt(s) < t), so that  !(sqrt(s2) < t)  <=>  !(s2 < s)  for every
// s2 >= 0 (sqrt is correctly rounded and monotonic).  Gesture statistics, :850-873.
double mag2_threshold(double t)
{
    if (!(t > 0.0)) return 0.0;                    // t <= 0 or NaN: nothing is below the threshold
    if (std::isinf(t)) return t;
    double s = t * t;
    if (std::isinf(s)) return s;
    while (std::sqrt(s) < t) s = std::nextafter(s, INFINITY);
    for (;;) {
        const double p = std::nextafter(s, 0.0);
        if (p == s || std::sqrt(p) < t) break;
        s = p;
    }
    return s;
}

// Brackets one kernel launch with events on the launching stream when timing is enabled.
struct TimedLaunch {
    of_ctx* ctx;
    cudaStream_t s;
    bool on;
    TimedLaunch(of_ctx* c, cudaStream_t st, int kind, int level) : ctx(c), s(st), on(c->timing)
    {
        // NVTX range per stage (the reference's per-stage timers, serial.cpp:204,261,303,320): free
        // when no profiler is attached
        char name[40];
        snprintf(name, sizeof name, kind == 0 ? "of_b200 pyrDown -> L%d" : "of_b200 lk_level L%d", level);
        nvtxRangePushA(name);
        if (!on) return;
        if (ctx->tcount >= 8192) { on = false; return; }
        if (ctx->tev.size() < 2 * (ctx->tcount + 1)) {
            cudaEvent_t e0 = nullptr, e1 = nullptr;
            if (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess) { on = false; return; }
            ctx->tev.push_back(e0);
            ctx->tev.push_back(e1);
            ctx->tkind.push_back(0);
            ctx->tlevel.push_back(0);
        }
        ctx->tkind[ctx->tcount] = kind;
        ctx->tlevel[ctx->tcount] = level;
        cudaEventRecord(ctx->tev[2 * ctx->tcount], s);
    }
    ~TimedLaunch()
    {
        nvtxRangePop();
        if (!on) return;
        cudaEventRecord(ctx->tev[2 * ctx->tcount + 1], s);
        ctx->tcount += 1;
    }
};

cudaError_t launch_lk_level(of_ctx* ctx, const ofb::LevelArgs& a, int n_pairs, cudaStream_t s)
{
    if (ctx->force_variant == 0 || ctx->force_variant == 4) {
        bool handled = false;
        cudaError_t e = ofb::launch_lk_level_phase(a, n_pairs, s, &ctx->launches, &handled);
        if (handled || e != cudaSuccess) return e;
    }
    return ofb::launch_lk_level_tile(a, n_pairs, s, &ctx->launches);
}

// The launch sequence of one batch on stream s: pyramids, then one fused kernel per level, coarse to fine.
of_status enqueue_batch(of_ctx* ctx, const Plan& p, int n_pairs, const double* img1, const double* img2, size_t in_pitch,
                        size_t in_stride, int L, double* u, double* v, size_t uv_pair_stride, size_t uv_step,
                        unsigned char* mask, size_t mask_pair_stride, size_t mask_step, of_gesture_stats* stats,
                        cudaStream_t s)
{
    char* base = ctx->scratch.base;
    const bool shared = p.shared;
    // ---- Gaussian pyramids (getGaussianPyramid :566-575, both frames :587-597) ------------
    for (int l = 1; l < L; ++l) {
        const LevelGeom& gs = p.g[l - 1];
        const LevelGeom& gd = p.g[l];
        double* dst = (double*)(base + p.pyr_off[l]);
        ofb::PyrDownArgs a{};
        a.w = gs.w;
        a.h = gs.h;
        a.dst_pitch = gd.pitch;
        a.dst_img_stride = gd.img_stride;
        if (l == 1) {
            a.src_pitch = in_pitch;
            a.src_img_stride = in_stride;
            if (shared) {
                a.src = img1; a.dst = dst; a.n_images = n_pairs + 1;
                TimedLaunch t(ctx, s, 0, l);
                OF_CUDA(ctx, ofb::launch_pyr_down(a, s, &ctx->launches));
            } else {
                a.src = img1; a.dst = dst; a.n_images = n_pairs;
                {
                    TimedLaunch t(ctx, s, 0, l);
                    OF_CUDA(ctx, ofb::launch_pyr_down(a, s, &ctx->launches));
                }
                a.src = img2; a.dst = dst + gd.img_stride * n_pairs;
                TimedLaunch t(ctx, s, 0, l);
                OF_CUDA(ctx, ofb::launch_pyr_down(a, s, &ctx->launches));
            }
        } else {
            a.src = (const double*)(base + p.pyr_off[l - 1]);
            a.src_pitch = gs.pitch;
            a.src_img_stride = gs.img_stride;
            a.dst = dst;
            a.n_images = p.n_frames;
            TimedLaunch t(ctx, s, 0, l);
            OF_CUDA(ctx, ofb::launch_pyr_down(a, s, &ctx->launches));
        }
    }

    if (stats) OF_CUDA(ctx, cudaMemsetAsync(stats, 0, sizeof(of_gesture_stats) * n_pairs, s));

    // ---- coarse to fine (:610-703), one fused launch per level ----------------------------
    for (int l = L - 1; l >= 0; --l) {
        const LevelGeom& g = p.g[l];
        ofb::LevelArgs a{};
        a.w = g.w;
        a.h = g.h;
        a.radius = ctx->params.win_radius;
        a.det_eps = ctx->params.det_eps;
        a.mag_thresh = ctx->params.mag_thresh;
        a.mag2_thresh = mag2_threshold(ctx->params.mag_thresh);
        if (l == 0) {
            a.A = img1;
            a.B = img2;
            a.img_pitch = in_pitch;
            a.a_pair_stride = a.b_pair_stride = in_stride;
        } else {
            const double* pyr = (const double*)(base + p.pyr_off[l]);
            a.A = pyr;
            a.B = shared ? pyr + g.img_stride : pyr + g.img_stride * n_pairs;
            a.img_pitch = g.pitch;
            a.a_pair_stride = a.b_pair_stride = g.img_stride;
        }
        if (l < L - 1) {
            const LevelGeom& gc = p.g[l + 1];
            a.cu = (const double*)(base + p.flow_u_off[l + 1]);
            a.cv = (const double*)(base + p.flow_v_off[l + 1]);
            a.c_pitch = gc.pitch;
            a.c_pair_stride = gc.img_stride;
            a.wc = gc.w;
            a.hc = gc.h;
        }
        if (l == 0) {
            if (u) {
                a.u = u;
                a.v = v;
                a.uv_pitch = uv_step / sizeof(double);
                a.uv_pair_stride = uv_pair_stride / sizeof(double);
            } else {                                   // statistics / mask only: scratch flow planes
                a.u = (double*)(base + p.flow_u_off[0]);
                a.v = (double*)(base + p.flow_v_off[0]);
                a.uv_pitch = g.pitch;
                a.uv_pair_stride = g.img_stride;
            }
            a.mask = mask;
            a.mask_pitch = mask_step;
            a.mask_pair_stride = mask_pair_stride;
            a.stats = stats;
        } else {
            a.u = (double*)(base + p.flow_u_off[l]);
            a.v = (double*)(base + p.flow_v_off[l]);
            a.uv_pitch = g.pitch;
            a.uv_pair_stride = g.img_stride;
        }
        TimedLaunch t(ctx, s, 1, l);
        OF_CUDA(ctx, launch_lk_level(ctx, a, n_pairs, s));
    }
    if (stats)
        OF_CUDA(ctx, ofb::launch_stats_finalize(stats, n_pairs, (double)ctx->params.count_thresh, s, &ctx->launches));
    return OF_OK;
}

// Core: all device pointers, stream-ordered.
of_status run_batch(of_ctx* ctx, int n_pairs, const double* img1, const double* img2,
                    size_t img_pair_stride, int w, int h, size_t step, int L, double* u,
                    double* v, size_t uv_pair_stride, size_t uv_step, unsigned char* mask,
                    size_t mask_pair_stride, size_t mask_step, of_gesture_stats* stats,
                    cudaStream_t s)
{
    of_status st = check_common(ctx, img1, img2, w, h, step, L);
    if (st) return st;
    if (n_pairs < 1) return fail(ctx, OF_E_BADARG, "n_pairs must be >= 1");
    if ((u == nullptr) != (v == nullptr)) return fail(ctx, OF_E_BADARG, "u and v must both be given or both be NULL");
    if (!u && !stats && !mask) return fail(ctx, OF_E_BADARG, "nothing to compute: u, v, mask and stats are all NULL");
    if (u && (uv_step % sizeof(double) || uv_step < (size_t)w * sizeof(double)))
        return fail(ctx, OF_E_BADARG, "uv_step_bytes %zu invalid for width %d", uv_step, w);
    if (mask && mask_step < (size_t)w) return fail(ctx, OF_E_BADARG, "mask_step_bytes %zu < width", mask_step);
    if (mask && n_pairs > 1 && mask_pair_stride < mask_step * (size_t)(h - 1) + (size_t)w)
        return fail(ctx, OF_E_BADARG, "mask_pair_stride_bytes %zu is smaller than one mask plane", mask_pair_stride);
    if (u && n_pairs > 1 && uv_pair_stride < uv_step * (size_t)(h - 1) + (size_t)w * sizeof(double))
        return fail(ctx, OF_E_BADARG, "uv_pair_stride_bytes %zu is smaller than one flow plane", uv_pair_stride);
    if (n_pairs > 1 && img_pair_stride % sizeof(double))
        return fail(ctx, OF_E_BADARG, "img_pair_stride_bytes must be a multiple of 8");
    if (u && n_pairs > 1 && uv_pair_stride % sizeof(double))
        return fail(ctx, OF_E_BADARG, "uv_pair_stride_bytes must be a multiple of 8");
    if (ctx->params.win_radius < 1 || ctx->params.win_radius > 15)
        return fail(ctx, OF_E_BADARG, "win_radius must be in 1..15");
    OF_CUDA(ctx, cudaSetDevice(ctx->device));
    // scratch hand-over between streams (see of_ctx::ev_scratch)
    if (ctx->scratch_busy && ctx->last_stream != s) OF_CUDA(ctx, cudaStreamWaitEvent(s, ctx->ev_scratch, 0));

    const bool shared = n_pairs > 1 && (const char*)img2 == (const char*)img1 + img_pair_stride;
    Plan p = make_plan(w, h, L, n_pairs, shared, u == nullptr);
    st = arena_reserve(ctx, ctx->scratch, p.total);
    if (st) return st;
    char* base = ctx->scratch.base;
    const size_t in_pitch = step / sizeof(double);
    const size_t in_stride = img_pair_stride / sizeof(double);
    ctx->last.assign(L, of_ctx::LastLevel{});
    for (int l = 0; l < L; ++l)
        ctx->last[l] = {p.g[l].w, p.g[l].h, p.g[l].pitch, p.g[l].img_stride, l ? p.pyr_off[l] : 0,
                        l ? p.flow_u_off[l] : 0, l ? p.flow_v_off[l] : 0};
    ctx->last_pairs = n_pairs;
    ctx->last_shared = shared;
    ctx->last_stream = s;

    // Small batches are launch-bound (KAT-7: 14 launches for 21 MB; the live loop calls one pair at a
    // time, :831-832): their launch sequence is captured into a CUDA graph the second time the same
    // call (geometry, parameters and pointers) is made, and replayed from then on.
    GraphKey key{};
    GraphEntry* ge = nullptr;
    if (graph_eligible(ctx, n_pairs, w, h)) {
        key = {n_pairs, w, h, L, ctx->params.win_radius, ctx->params.count_thresh, ctx->params.det_eps, ctx->params.mag_thresh,
               img1, img2, u, v, mask, stats, img_pair_stride, step, uv_pair_stride, uv_step, mask_pair_stride, mask_step,
               ctx->scratch.base};
        ge = graph_lookup(ctx, key);
    }
    if (ge && ge->exec) {
        OF_CUDA(ctx, cudaGraphLaunch(ge->exec, s));
        ctx->launches += ge->launches;
        ctx->graph_replays += 1;
    } else if (ge && ge->seen) {
        // second sighting: capture on the context's own stream (a legacy default stream cannot capture)
        const unsigned long long l0 = ctx->launches;
        OF_CUDA(ctx, cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeRelaxed));
        st = enqueue_batch(ctx, p, n_pairs, img1, img2, in_pitch, in_stride, L, u, v, uv_pair_stride, uv_step, mask,
                           mask_pair_stride, mask_step, stats, ctx->stream);
        cudaGraph_t g = nullptr;
        cudaError_t e = cudaStreamEndCapture(ctx->stream, &g);
        if (st) {
            if (g) cudaGraphDestroy(g);
            return st;
        }
        if (e == cudaSuccess) e = cudaGraphInstantiate(&ge->exec, g, 0);
        if (g) cudaGraphDestroy(g);
        if (e != cudaSuccess) {
            ge->exec = nullptr;
            return fail(ctx, OF_E_CUDA, "CUDA graph capture failed: %s", cudaGetErrorString(e));
        }
        ge->launches = ctx->launches - l0;
        OF_CUDA(ctx, cudaGraphLaunch(ge->exec, s));
        ctx->graph_replays += 1;
    } else {
        if (ge) ge->seen = true;
        st = enqueue_batch(ctx, p, n_pairs, img1, img2, in_pitch, in_stride, L, u, v, uv_pair_stride, uv_step, mask,
                           mask_pair_stride, mask_step, stats, s);
        if (st) return st;
    }
    OF_CUDA(ctx, cudaEventRecord(ctx->ev_scratch, s));
    ctx->scratch_busy = true;
    return OF_OK;
}

// The device-pointer, batch and stream entry points read their inputs through const pointers and cannot
// leave the warped frames in them (quirk Q1); only the host-pointer of_ctf_f64 honours the flag.
of_status reject_write_back(of_ctx* ctx, int n_levels, const char* who)
{
    if (ctx->params.write_back_warped && n_levels > 1)
        return fail(ctx, OF_E_BADARG, "%s cannot honour write_back_warped = 1 (inputs are const); use of_ctf_f64", who);
    return OF_OK;
}

cudaStream_t pick(of_ctx* ctx, void* stream) { return stream ? (cudaStream_t)stream : ctx->stream; }

}  // namespace

extern "C" {

void of_params_default(of_params* p)
{
    if (!p) return;
    p->win_radius = 1;
    p->det_eps = 0.0;
    p->write_back_warped = 0;
    p->mag_thresh = 20.0;
    p->count_thresh = 500;
}

int of_abi_version(void) { return OF_B200_ABI_VERSION; }

int of_get_max_layer(int width, int height)
{
    if (width <= 0 || height <= 0) return 0;
    int pw = 1, ph = 1;
    while (pw < 31 && width % (1 << pw) == 0) ++pw;
    while (ph < 31 && height % (1 << ph) == 0) ++ph;
    return pw < ph ? pw : ph;
}

of_status of_ctx_create(int device, const of_params* params, of_ctx** out)
{
    if (!out) return fail(nullptr, OF_E_BADARG, "out is NULL");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail(nullptr, OF_E_CUDA, "no usable CUDA device (%s); this library has no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    }
    if (device < 0 || device >= n) return fail(nullptr, OF_E_BADARG, "device %d out of range (0..%d)", device, n - 1);
    cudaDeviceProp prop{};
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) return fail(nullptr, OF_E_CUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(e));
    if (prop.major != 10)
        return fail(nullptr, OF_E_CUDA, "device %d is sm_%d%d; this library is built for sm_100a (B200) only",
                    device, prop.major, prop.minor);
    of_ctx* ctx = new (std::nothrow) of_ctx();
    if (!ctx) return fail(nullptr, OF_E_NOMEM, "out of host memory");
    ctx->device = device;
    of_params_default(&ctx->params);
    if (params) {
        if (const char* why = validate_params(*params)) {
            delete ctx;
            return fail(nullptr, OF_E_BADARG, "%s", why);
        }
        ctx->params = *params;
    }
    if (const char* gm = getenv("OF_B200_GRAPH")) ctx->graph_mode = atoi(gm);
    if (const char* fv = getenv("OF_B200_VARIANT")) {
        if (!strcmp(fv, "tile")) ctx->force_variant = 1;
        else if (!strcmp(fv, "phase")) ctx->force_variant = 4;
    }
    e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->down_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->ev_scratch, cudaEventDisableTiming);
    for (int i = 0; i < 2 && e == cudaSuccess; ++i) {
        e = cudaEventCreateWithFlags(&ctx->ev_up[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->ev_done[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->ev_down[i], cudaEventDisableTiming);
    }
    if (e != cudaSuccess) {
        of_status st = fail(nullptr, OF_E_CUDA, "context setup failed: %s", cudaGetErrorString(e));
        of_ctx_destroy(ctx);
        return st;
    }
    *out = ctx;
    return OF_OK;
}

void of_ctx_destroy(of_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    graph_clear(ctx);
    if (ctx->live.base) cudaFree(ctx->live.base);
    if (ctx->rz.xofs) cudaFree(ctx->rz.xofs);
    if (ctx->scratch.base) cudaFree(ctx->scratch.base);
    if (ctx->staging.base) cudaFree(ctx->staging.base);
    if (ctx->h_stats) cudaFreeHost(ctx->h_stats);
    for (cudaEvent_t e : ctx->tev) cudaEventDestroy(e);
    if (ctx->ev_scratch) cudaEventDestroy(ctx->ev_scratch);
    for (int i = 0; i < 2; ++i) {
        if (ctx->ev_up[i]) cudaEventDestroy(ctx->ev_up[i]);
        if (ctx->ev_done[i]) cudaEventDestroy(ctx->ev_done[i]);
        if (ctx->ev_down[i]) cudaEventDestroy(ctx->ev_down[i]);
    }
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->down_stream) cudaStreamDestroy(ctx->down_stream);
    delete ctx;
}

of_status of_ctx_set_params(of_ctx* ctx, const of_params* params)
{
    if (!ctx || !params) return OF_E_BADARG;
    if (const char* why = validate_params(*params)) return fail(ctx, OF_E_BADARG, "%s", why);
    ctx->params = *params;
    return OF_OK;
}

const char* of_last_error_string(const of_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

unsigned long long of_ctx_launch_count(const of_ctx* ctx) { return ctx ? ctx->launches : 0ull; }

unsigned long long of_ctx_graph_replay_count(const of_ctx* ctx) { return ctx ? ctx->graph_replays : 0ull; }

of_status of_ctx_set_timing(of_ctx* ctx, int enable)
{
    if (!ctx) return OF_E_BADARG;
    ctx->timing = enable != 0;
    ctx->tcount = 0;
    return OF_OK;
}

of_status of_ctx_get_timing(of_ctx* ctx, int max_entries, int* kinds, int* levels, float* ms, int* n)
{
    if (!ctx || !n || (max_entries > 0 && (!kinds || !levels || !ms))) return OF_E_BADARG;
    OF_CUDA(ctx, cudaSetDevice(ctx->device));
    int cnt = 0;
    for (size_t i = 0; i < ctx->tcount && cnt < max_entries; ++i, ++cnt) {
        OF_CUDA(ctx, cudaEventSynchronize(ctx->tev[2 * i + 1]));
        OF_CUDA(ctx, cudaEventElapsedTime(&ms[cnt], ctx->tev[2 * i], ctx->tev[2 * i + 1]));
        kinds[cnt] = ctx->tkind[i];
        levels[cnt] = ctx->tlevel[i];
    }
    *n = cnt;
    return OF_OK;
}

of_status of_ctx_get_level_f64(of_ctx* ctx, int what, int level, int pair, double* dst,
                               size_t dst_step_bytes, int* width, int* height)
{
    if (!ctx) return OF_E_BADARG;
    if (ctx->last.empty()) return fail(ctx, OF_E_BADARG, "no pyramid call has run on this context yet");
    if (level < 1 || level >= (int)ctx->last.size())
        return fail(ctx, OF_E_BADARG, "level %d: only levels 1..%d live in the context (level 0 is the caller's)", level,
                    (int)ctx->last.size() - 1);
    if (pair < 0 || pair >= ctx->last_pairs) return fail(ctx, OF_E_BADARG, "pair %d out of range", pair);
    if (what < OF_LEVEL_FRAME1 || what > OF_LEVEL_FLOW_V) return fail(ctx, OF_E_BADARG, "unknown plane %d", what);
    const of_ctx::LastLevel& g = ctx->last[level];
    if (width) *width = g.w;
    if (height) *height = g.h;
    if (!dst) return OF_OK;                                  // size query
    if (dst_step_bytes < (size_t)g.w * sizeof(double)) return fail(ctx, OF_E_BADARG, "dst_step_bytes too small");
    const char* base = ctx->scratch.base;
    const double* src;
    if (what == OF_LEVEL_FRAME1)
        src = (const double*)(base + g.pyr_off) + g.img_stride * pair;
    else if (what == OF_LEVEL_FRAME2)
        src = (const double*)(base + g.pyr_off) + g.img_stride * (ctx->last_shared ? pair + 1 : ctx->last_pairs + pair);
    else
        src = (const double*)(base + (what == OF_LEVEL_FLOW_U ? g.flow_u_off : g.flow_v_off)) + g.img_stride * pair;
    OF_CUDA(ctx, cudaSetDevice(ctx->device));
    OF_CUDA(ctx, cudaEventSynchronize(ctx->ev_scratch));     // not the stream: the caller may have destroyed it
    OF_CUDA(ctx, cudaMemcpy2D(dst, dst_step_bytes, src, g.pitch * sizeof(double), (size_t)g.w * sizeof(double), g.h,
                              cudaMemcpyDefault));
    return OF_OK;
}

of_status of_ctx_synchronize(of_ctx* ctx)
{
    if (!ctx) return OF_E_BADARG;
    OF_CUDA(ctx, cudaSetDevice(ctx->device));
    OF_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    OF_CUDA(ctx, cudaStreamSynchronize(ctx->copy_stream));
    OF_CUDA(ctx, cudaStreamSynchronize(ctx->down_stream));
    return OF_OK;
}

of_status of_ctf_f64_batch_dev(of_ctx* ctx, int n_pairs, const double* img1, const double* img2,
                               size_t img_pair_stride_bytes, int width, int height,
                               size_t step_bytes, int n_levels, double* u, double* v,
                               size_t uv_pair_stride_bytes, size_t uv_step_bytes,
                               unsigned char* mask, size_t mask_pair_stride_bytes,
                               size_t mask_step_bytes, of_gesture_stats* stats, void* stream)
{
    if (!ctx) return OF_E_BADARG;
    if (of_status st = reject_write_back(ctx, n_levels, "of_ctf_f64_batch_dev")) return st;
    return run_batch(ctx, n_pairs, img1, img2, img_pair_stride_bytes, width, height, step_bytes,
                     n_levels, u, v, uv_pair_stride_bytes, uv_step_bytes, mask,
                     mask_pair_stride_bytes, mask_step_bytes, stats, pick(ctx, stream));
}

of_status of_ctf_f64_dev(of_ctx* ctx, const double* img1, const double* img2, int width, int height,
                         size_t step_bytes, int n_levels, double* u, double* v,
                         size_t uv_step_bytes, unsigned char* mask, size_t mask_step_bytes,
                         void* stream)
{
    if (!ctx) return OF_E_BADARG;
    if (!u || !v) return fail(ctx, OF_E_BADARG, "null flow pointer");
    if (of_status st = reject_write_back(ctx, n_levels, "of_ctf_f64_dev")) return st;
    return run_batch(ctx, 1, img1, img2, 0, width, height, step_bytes, n_levels, u, v, 0,
                     uv_step_bytes, mask, 0, mask_step_bytes, nullptr, pick(ctx, stream));
}

of_status of_lk_single_f64_dev(of_ctx* ctx, const double* img1, const double* img2, int width,
                               int height, size_t step_bytes, double* u, double* v,
                               size_t uv_step_bytes, unsigned char* mask, size_t mask_step_bytes,
                               void* stream)
{
    return of_ctf_f64_dev(ctx, img1, img2, width, height, step_bytes, 1, u, v, uv_step_bytes, mask,
                          mask_step_bytes, stream);
}

of_status of_gesture_stats_dev(of_ctx* ctx, int n_pairs, const double* u, const double* v,
                               size_t uv_pair_stride_bytes, int width, int height,
                               size_t uv_step_bytes, of_gesture_stats* stats, void* stream)
{
    if (!ctx) return OF_E_BADARG;
    if (!u || !v || !stats || n_pairs < 1 || width < 1 || height < 1)
        return fail(ctx, OF_E_BADARG, "bad argument to of_gesture_stats_dev");
    if (uv_step_bytes % 8 || uv_step_bytes < (size_t)width * 8 || uv_pair_stride_bytes % 8)
        return fail(ctx, OF_E_BADARG, "bad step/stride");
    OF_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = pick(ctx, stream);
    OF_CUDA(ctx, cudaMemsetAsync(stats, 0, sizeof(of_gesture_stats) * n_pairs, s));
    ofb::StatsArgs a{};
    a.u = u;
    a.v = v;
    a.uv_pitch = uv_step_bytes / 8;
    a.uv_pair_stride = uv_pair_stride_bytes / 8;
    a.w = width;
    a.h = height;
    a.stats = stats;
    a.mag_thresh = ctx->params.mag_thresh;
    OF_CUDA(ctx, ofb::launch_gesture_stats(a, n_pairs, s, &ctx->launches));
    OF_CUDA(ctx, ofb::launch_stats_finalize(stats, n_pairs, (double)ctx->params.count_thresh, s, &ctx->launches));
    return OF_OK;
}

of_status of_pyr_down_f64_dev(of_ctx* ctx, const double* src, int width, int height,
                              size_t step_bytes, double* dst, size_t dst_step_bytes, void* stream)
{
    if (!ctx) return OF_E_BADARG;
    if (!src || !dst || width < 1 || height < 1) return fail(ctx, OF_E_BADARG, "bad argument to of_pyr_down_f64_dev");
    if (step_bytes % 8 || step_bytes < (size_t)width * 8 || dst_step_bytes % 8 ||
        dst_step_bytes < (size_t)((width + 1) / 2) * 8)
        return fail(ctx, OF_E_BADARG, "bad step");
    OF_CUDA(ctx, cudaSetDevice(ctx->device));
    ofb::PyrDownArgs a{};
    a.src = src;
    a.src_pitch = step_bytes / 8;
    a.dst = dst;
    a.dst_pitch = dst_step_bytes / 8;
    a.w = width;
    a.h = height;
    a.n_images = 1;
    OF_CUDA(ctx, ofb::launch_pyr_down(a, pick(ctx, stream), &ctx->launches));
    return OF_OK;
}

/* ---- live-loop pre-processing (ParallelVideoPyr.cpp:791-813) ------------------------------ */

of_status of_bgr2gray_u8_dev(of_ctx* ctx, const unsigned char* bgr, int width, int height,
                             size_t bgr_step_bytes, unsigned char* gray, size_t gray_step_bytes,
                             int coeff_bits, void* stream)
{
    if (!ctx) return OF_E_BADARG;
    if (!bgr || !gray || width < 1 || height < 1) return fail(ctx, OF_E_BADARG, "bad argument to of_bgr2gray_u8_dev");
    if (bgr_step_bytes < (size_t)width * 3 || gray_step_bytes < (size_t)width) return fail(ctx, OF_E_BADARG, "bad step");
    if (coeff_bits != 14 && coeff_bits != 15) return fail(ctx, OF_E_BADARG, "coeff_bits must be 14 (OpenCV 2.4) or 15 (OpenCV >= 3)");
    OF_CUDA(ctx, cudaSetDevice(ctx->device));
    OF_CUDA(ctx, ofb::launch_bgr2gray(bgr, bgr_step_bytes, gray, gray_step_bytes, width, height, coeff_bits,
                                      pick(ctx, stream), &ctx->launches));
    return OF_OK;
}

of_status of_live_diff_f64_dev(of_ctx* ctx, const unsigned char* cur_gray, const unsigned char* prev_gray,
                               int width, int height, size_t gray_step_bytes, double learning_rate,
                               int diff_thresh, int morph_iterations, double* out, size_t out_step_bytes,
                               void* stream)
{
    if (!ctx) return OF_E_BADARG;
    if (!cur_gray || !prev_gray || !out || width < 1 || height < 1)
        return fail(ctx, OF_E_BADARG, "bad argument to of_live_diff_f64_dev");
    if (gray_step_bytes < (size_t)width || out_step_bytes % 8 || out_step_bytes < (size_t)width * 8)
        return fail(ctx, OF_E_BADARG, "bad step");
    if (morph_iterations < 1 || morph_iterations > 4) return fail(ctx, OF_E_BADARG, "morph_iterations must be in 1..4");
    if (!(learning_rate == learning_rate)) return fail(ctx, OF_E_BADARG, "learning_rate is NaN");
    OF_CUDA(ctx, cudaSetDevice(ctx->device));
    ofb::LiveDiffArgs a{};
    a.cur = cur_gray;
    a.prev = prev_gray;
    a.gray_step = gray_step_bytes;
    a.w = width;
    a.h = height;
    a.beta = (float)(-learning_rate);
    a.thresh = diff_thresh;
    a.iterations = morph_iterations;
    a.out = out;
    a.out_pitch = out_step_bytes / 8;
    OF_CUDA(ctx, ofb::launch_live_diff(a, pick(ctx, stream), &ctx->launches));
    return OF_OK;
}

of_status of_live_diff_f64(of_ctx* ctx, const unsigned char* cur, const unsigned char* prev, int channels,
                           int width, int height, size_t frame_step_bytes, int coeff_bits,
                           double learning_rate, int diff_thresh, int morph_iterations, double* out,
                           size_t out_step_bytes)
{
    if (!ctx) return OF_E_BADARG;
    if (!cur || !prev || !out || width < 1 || height < 1) return fail(ctx, OF_E_BADARG, "bad argument to of_live_diff_f64");
    if (channels != 1 && channels != 3) return fail(ctx, OF_E_BADARG, "channels must be 1 (gray) or 3 (BGR)");
    if (frame_step_bytes < (size_t)width * channels) return fail(ctx, OF_E_BADARG, "bad frame_step_bytes");
    if (out_step_bytes % 8 || out_step_bytes < (size_t)width * 8) return fail(ctx, OF_E_BADARG, "bad out_step_bytes");
    OF_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t in_pitch = align_up((size_t)width * channels, 256);
    const size_t g_pitch = align_up((size_t)width, 256);
    const size_t o_pitch = align_up((size_t)width, 16);                  // doubles
    const size_t in_plane = in_pitch * height, g_plane = g_pitch * height;
    const size_t need = 2 * in_plane + 2 * g_plane + o_pitch * height * 8 + 1024;
    of_status st = arena_reserve(ctx, ctx->staging, need);
    if (st) return st;
    unsigned char* d_in = (unsigned char*)ctx->staging.base;
    unsigned char* d_g = d_in + 2 * in_plane;
    double* d_out = (double*)(ctx->staging.base + align_up(2 * in_plane + 2 * g_plane, 256));
    cudaStream_t s = ctx->stream;
    const unsigned char* src[2] = {cur, prev};
    const unsigned char* gray[2];
    for (int k = 0; k < 2; ++k) {
        OF_CUDA(ctx, cudaMemcpy2DAsync(d_in + k * in_plane, in_pitch, src[k], frame_step_bytes, (size_t)width * channels,
                                       height, cudaMemcpyHostToDevice, s));
        if (channels == 3) {
            st = of_bgr2gray_u8_dev(ctx, d_in + k * in_plane, width, height, in_pitch, d_g + k * g_plane, g_pitch,
                                    coeff_bits, s);
            if (st) return st;
            gray[k] = d_g + k * g_plane;
        } else {
            gray[k] = d_in + k * in_plane;
        }
    }
    st = of_live_diff_f64_dev(ctx, gray[0], gray[1], width, height, channels == 3 ? g_pitch : in_pitch, learning_rate,
                              diff_thresh, morph_iterations, d_out, o_pitch * 8, s);
    if (st) return st;
    OF_CUDA(ctx, cudaMemcpy2DAsync(out, out_step_bytes, d_out, o_pitch * 8, (size_t)width * 8, height,
                                   cudaMemcpyDeviceToHost, s));
    OF_CUDA(ctx, cudaStreamSynchronize(s));
    return OF_OK;
}

}  // extern "C"

/* ---- live-loop colour-frame steps and the whole iteration (ParallelVideoPyr.cpp:769-905) ---------- */

namespace {

// Device-resident bicubic tables for (sw, sh) -> (dw, dh), built once per geometry.
of_status resize_tables(of_ctx* ctx, int sw, int sh, int dw, int dh)
{
    of_ctx::ResizeTab& t = ctx->rz;
    if (t.xofs && t.sw == sw && t.sh == sh && t.dw == dw && t.dh == dh) return OF_OK;
    if (t.xofs) {
        OF_CUDA(ctx, cudaDeviceSynchronize());
        cudaFree(t.xofs);
        t = of_ctx::ResizeTab{};
    }
    std::vector<int> xo(4 * (size_t)dw), yo(4 * (size_t)dh);
    std::vector<short> xc(4 * (size_t)dw), yc(4 * (size_t)dh);
    std::vector<float> yb(4 * (size_t)dh);
    ofb::build_cubic_table(sw, dw, xo.data(), xc.data());
    ofb::build_cubic_table(sh, dh, yo.data(), yc.data());
    const float scale = 1.0f / (2048.0f * 2048.0f);
    for (size_t i = 0; i < yb.size(); ++i) yb[i] = (float)yc[i] * scale;
    const size_t b_xo = align_up(xo.size() * 4, 256), b_yo = align_up(yo.size() * 4, 256);
    const size_t b_yb = align_up(yb.size() * 4, 256), b_xc = align_up(xc.size() * 2, 256);
    char* base = nullptr;
    cudaError_t e = cudaMalloc((void**)&base, b_xo + b_yo + b_yb + b_xc);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(ctx, OF_E_NOMEM, "cudaMalloc of the resize tables failed: %s", cudaGetErrorString(e));
    }
    t.xofs = (int*)base;
    t.yofs = (int*)(base + b_xo);
    t.ybeta = (float*)(base + b_xo + b_yo);
    t.xco = (short*)(base + b_xo + b_yo + b_yb);
    OF_CUDA(ctx, cudaMemcpy(t.xofs, xo.data(), xo.size() * 4, cudaMemcpyHostToDevice));
    OF_CUDA(ctx, cudaMemcpy(t.yofs, yo.data(), yo.size() * 4, cudaMemcpyHostToDevice));
    OF_CUDA(ctx, cudaMemcpy(t.ybeta, yb.data(), yb.size() * 4, cudaMemcpyHostToDevice));
    OF_CUDA(ctx, cudaMemcpy(t.xco, xc.data(), xc.size() * 2, cudaMemcpyHostToDevice));
    t.sw = sw; t.sh = sh; t.dw = dw; t.dh = dh;
    return OF_OK;
}

}  // namespace

extern "C" {

of_status of_resize_cubic_u8_dev(of_ctx* ctx, const unsigned char* src, int src_width, int src_height,
                                 size_t src_step_bytes, int channels, unsigned char* dst, int dst_width,
                                 int dst_height, size_t dst_step_bytes, void* stream)
{
    if (!ctx) return OF_E_BADARG;
    if (!src || !dst || src_width < 1 || src_height < 1 || dst_width < 1 || dst_height < 1)
        return fail(ctx, OF_E_BADARG, "bad argument to of_resize_cubic_u8_dev");
    if (channels != 1 && channels != 3) return fail(ctx, OF_E_BADARG, "channels must be 1 or 3");
    if (src_step_bytes < (size_t)src_width * channels || dst_step_bytes < (size_t)dst_width * channels)
        return fail(ctx, OF_E_BADARG, "bad step");
    OF_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = pick(ctx, stream);
    if (src_width == dst_width && src_height == dst_height) {          // cv::resize copies when the sizes agree
        OF_CUDA(ctx, cudaMemcpy2DAsync(dst, dst_step_bytes, src, src_step_bytes, (size_t)src_width * channels, src_height,
                                       cudaMemcpyDeviceToDevice, s));
        return OF_OK;
    }
    if (of_status st = resize_tables(ctx, src_width, src_height, dst_width, dst_height)) return st;
    OF_CUDA(ctx, ofb::launch_resize_cubic(src, src_step_bytes, channels, dst, dst_step_bytes, dst_width, dst_height,
                                          ctx->rz.xofs, ctx->rz.xco, ctx->rz.yofs, ctx->rz.ybeta, s, &ctx->launches));
    return OF_OK;
}

of_status of_gaussian_blur_u8_dev(of_ctx* ctx, const unsigned char* src, int width, int height, size_t src_step_bytes,
                                  int channels, int ksize, double sigma, int gray_coeff_bits, unsigned char* dst,
                                  size_t dst_step_bytes, void* stream)
{
    if (!ctx) return OF_E_BADARG;
    if (!src || !dst || width < 1 || height < 1) return fail(ctx, OF_E_BADARG, "bad argument to of_gaussian_blur_u8_dev");
    if (channels != 1 && channels != 3) return fail(ctx, OF_E_BADARG, "channels must be 1 or 3");
    if (gray_coeff_bits != 0 && gray_coeff_bits != 14 && gray_coeff_bits != 15)
        return fail(ctx, OF_E_BADARG, "gray_coeff_bits must be 0, 14 or 15");
    if (gray_coeff_bits && channels != 3) return fail(ctx, OF_E_BADARG, "the fused cvtColor needs channels = 3");
    if (src_step_bytes < (size_t)width * channels || dst_step_bytes < (size_t)width * (gray_coeff_bits ? 1 : channels))
        return fail(ctx, OF_E_BADARG, "bad step");
    int k[31];
    if (!ofb::build_gauss_kernel_q8(ksize, sigma, k)) return fail(ctx, OF_E_BADARG, "ksize must be odd, 1..31, sigma > 0");
    OF_CUDA(ctx, cudaSetDevice(ctx->device));
    OF_CUDA(ctx, ofb::launch_gauss_u8(src, src_step_bytes, channels, dst, dst_step_bytes, width, height, k, ksize,
                                      gray_coeff_bits, pick(ctx, stream), &ctx->launches));
    return OF_OK;
}

void of_live_params_default(of_live_params* p)
{
    if (!p) return;
    p->work_width = p->work_height = 1080;
    p->blur_ksize = 9;
    p->blur_sigma = 1.5;
    p->coeff_bits = 14;
    p->learning_rate = 0.3;
    p->diff_thresh = 10;
    p->morph_iterations = 2;
    p->n_levels = 0;
}

of_status of_live_reset(of_ctx* ctx)
{
    if (!ctx) return OF_E_BADARG;
    ctx->ls = of_ctx::LiveState{};
    return OF_OK;
}

of_status of_live_step_u8(of_ctx* ctx, const of_live_params* lp_in, const unsigned char* frame_bgr, int width, int height,
                          size_t step_bytes, of_gesture_stats* stats, double* u, double* v, size_t uv_step_bytes,
                          int* stage)
{
    if (!ctx) return OF_E_BADARG;
    of_live_params lp;
    of_live_params_default(&lp);
    if (lp_in) lp = *lp_in;
    if (!frame_bgr || width < 1 || height < 1 || step_bytes < (size_t)width * 3)
        return fail(ctx, OF_E_BADARG, "bad frame argument to of_live_step_u8");
    if (!stage) return fail(ctx, OF_E_BADARG, "stage is NULL");
    if ((u == nullptr) != (v == nullptr)) return fail(ctx, OF_E_BADARG, "u and v must both be given or both be NULL");
    const int W = lp.work_width, H = lp.work_height;
    if (W < 1 || H < 1 || W > 32767 || H > 32767) return fail(ctx, OF_E_BADARG, "bad working size");
    if (u && (uv_step_bytes % 8 || uv_step_bytes < (size_t)W * 8)) return fail(ctx, OF_E_BADARG, "bad uv_step_bytes");
    if (lp.coeff_bits != 14 && lp.coeff_bits != 15) return fail(ctx, OF_E_BADARG, "coeff_bits must be 14 or 15");
    if (lp.morph_iterations < 1 || lp.morph_iterations > 4) return fail(ctx, OF_E_BADARG, "morph_iterations must be in 1..4");
    if (!(lp.learning_rate == lp.learning_rate)) return fail(ctx, OF_E_BADARG, "learning_rate is NaN");
    int k[31];
    if (!ofb::build_gauss_kernel_q8(lp.blur_ksize, lp.blur_sigma, k)) return fail(ctx, OF_E_BADARG, "bad blur kernel");
    const int L = lp.n_levels > 0 ? lp.n_levels : of_get_max_layer(W, H);
    if (of_status st = check_common(ctx, frame_bgr, frame_bgr, W, H, (size_t)W * 8, L)) return st;
    OF_CUDA(ctx, cudaSetDevice(ctx->device));

    // device buffers: captured frame, resized frame, two gray frames, two difference images, flow planes
    const size_t f_pitch = align_up((size_t)width * 3, 256), r_pitch = align_up((size_t)W * 3, 256);
    const size_t g_pitch = align_up((size_t)W, 256), d_pitch = align_up((size_t)W, 16);          // d_pitch in doubles
    const size_t b_frame = f_pitch * height, b_res = r_pitch * H, b_gray = g_pitch * H;
    const size_t b_diff = align_up(d_pitch * H * 8, 256);
    const size_t need = b_frame + b_res + 2 * b_gray + 4 * b_diff + 256;
    if (need > ctx->live.cap) ctx->ls = of_ctx::LiveState{};                 // the state lives in the arena
    if (ctx->ls.w != W || ctx->ls.h != H) ctx->ls = of_ctx::LiveState{};
    if (of_status st = arena_reserve(ctx, ctx->live, need)) return st;
    unsigned char* d_frame = (unsigned char*)ctx->live.base;
    unsigned char* d_res = d_frame + b_frame;
    unsigned char* d_gray[2] = {d_res + b_res, d_res + b_res + b_gray};
    double* d_diff[2] = {(double*)(d_res + b_res + 2 * b_gray), (double*)(d_res + b_res + 2 * b_gray + b_diff)};
    double* d_u = (double*)(d_res + b_res + 2 * b_gray + 2 * b_diff);
    double* d_v = (double*)(d_res + b_res + 2 * b_gray + 3 * b_diff);
    of_ctx::LiveState& ls = ctx->ls;
    ls.w = W;
    ls.h = H;
    cudaStream_t s = ctx->stream;

    // :781-791  capture >> frame; resize; GaussianBlur; cvtColor
    OF_CUDA(ctx, cudaMemcpy2DAsync(d_frame, f_pitch, frame_bgr, step_bytes, (size_t)width * 3, height, cudaMemcpyHostToDevice, s));
    const unsigned char* blur_src = d_frame;
    size_t blur_pitch = f_pitch;
    if (width != W || height != H) {
        if (of_status st = of_resize_cubic_u8_dev(ctx, d_frame, width, height, f_pitch, 3, d_res, W, H, r_pitch, s)) return st;
        blur_src = d_res;
        blur_pitch = r_pitch;
    }
    const int g_cur = ls.frames == 0 ? 0 : 1 - ls.cur_gray;
    OF_CUDA(ctx, ofb::launch_gauss_u8(blur_src, blur_pitch, 3, d_gray[g_cur], g_pitch, W, H, k, lp.blur_ksize, lp.coeff_bits, s,
                                      &ctx->launches));
    if (ls.frames == 0) {                                  // :793-797 first frame
        ls.cur_gray = g_cur;
        ls.frames = 1;
        *stage = 0;
        OF_CUDA(ctx, cudaStreamSynchronize(s));            // the caller may reuse its frame buffer
        return OF_OK;
    }
    // :800-813  diff = cur - 0.3 prev; threshold; Sobel x + y; dilate; erode
    const int d_cur = ls.frames == 1 ? 0 : 1 - ls.cur_diff;
    if (of_status st = of_live_diff_f64_dev(ctx, d_gray[g_cur], d_gray[ls.cur_gray], W, H, g_pitch, lp.learning_rate,
                                            lp.diff_thresh, lp.morph_iterations, d_diff[d_cur], d_pitch * 8, s))
        return st;
    ls.cur_gray = g_cur;                                   // prevFrame = current_frame :801
    if (ls.frames == 1) {                                  // :816-820 first difference image
        ls.cur_diff = d_cur;
        ls.frames = 2;
        *stage = 1;
        OF_CUDA(ctx, cudaStreamSynchronize(s));
        return OF_OK;
    }
    // :831-832  coarseToFineEstimation(prevDiff, diff, u, v, getMaxLayer(prevDiff)) and :845-888 statistics
    const size_t st_off = 0;
    if (of_status st = arena_reserve(ctx, ctx->staging, 256)) return st;
    of_gesture_stats* d_stats = (of_gesture_stats*)(ctx->staging.base + st_off);
    of_status st = run_batch(ctx, 1, d_diff[ls.cur_diff], d_diff[d_cur], 0, W, H, d_pitch * 8, L, u ? d_u : nullptr,
                             u ? d_v : nullptr, 0, d_pitch * 8, nullptr, 0, 0, stats ? d_stats : nullptr, s);
    if (st) return st;
    if (!u && !stats) return fail(ctx, OF_E_BADARG, "nothing to compute: stats, u and v are all NULL");
    if (stats) OF_CUDA(ctx, cudaMemcpyAsync(stats, d_stats, sizeof(of_gesture_stats), cudaMemcpyDeviceToHost, s));
    if (u) {
        OF_CUDA(ctx, cudaMemcpy2DAsync(u, uv_step_bytes, d_u, d_pitch * 8, (size_t)W * 8, H, cudaMemcpyDeviceToHost, s));
        OF_CUDA(ctx, cudaMemcpy2DAsync(v, uv_step_bytes, d_v, d_pitch * 8, (size_t)W * 8, H, cudaMemcpyDeviceToHost, s));
    }
    int next_prev = d_cur;
    if (ctx->params.write_back_warped && L > 1) {
        // quirk Q1: the reference's diff has been overwritten by its warped version (:700-701) before
        // prevDiff = diff (:843): frame 2 sampled at -flow / 2 of level 1, into the other slot
        Plan p = make_plan(W, H, L, 1, false, u == nullptr);
        ofb::WarpArgs a{};
        a.cu = (const double*)(ctx->scratch.base + p.flow_u_off[1]);
        a.cv = (const double*)(ctx->scratch.base + p.flow_v_off[1]);
        a.c_pitch = p.g[1].pitch;
        a.wc = p.g[1].w;
        a.hc = p.g[1].h;
        a.w = W;
        a.h = H;
        a.src = d_diff[d_cur];
        a.src_pitch = d_pitch;
        a.dst = d_diff[ls.cur_diff];
        a.dst_pitch = d_pitch;
        a.sign = -1.0f;
        OF_CUDA(ctx, ofb::launch_warp_level(a, s, &ctx->launches));
        next_prev = ls.cur_diff;
    }
    ls.cur_diff = next_prev;                               // prevDiff = diff :843
    *stage = 2;
    OF_CUDA(ctx, cudaStreamSynchronize(s));
    return OF_OK;
}

}  // extern "C"

extern "C" {

/* ---- host-pointer entry points ------------------------------------------------------- */

of_status of_ctf_f64(of_ctx* ctx, double* img1, double* img2, int width, int height,
                     size_t step_bytes, int n_levels, double* u, double* v, size_t uv_step_bytes,
                     unsigned char* mask, size_t mask_step_bytes)
{
    of_status st = check_common(ctx, img1, img2, width, height, step_bytes, n_levels);
    if (st) return st;
    if (!u || !v) return fail(ctx, OF_E_BADARG, "null flow pointer");
    if (uv_step_bytes % 8 || uv_step_bytes < (size_t)width * 8) return fail(ctx, OF_E_BADARG, "bad uv_step_bytes");
    if (mask && mask_step_bytes < (size_t)width) return fail(ctx, OF_E_BADARG, "mask_step_bytes < width");
    OF_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t pitch = align_up((size_t)width, 16);            // doubles
    const size_t plane = align_up(pitch * height * sizeof(double), 256);
    const size_t mplane = align_up(pitch * height, 256);
    st = arena_reserve(ctx, ctx->staging, 5 * plane + mplane);
    if (st) return st;
    char* b = ctx->staging.base;
    double* d1 = (double*)b;
    double* d2 = (double*)(b + plane);
    double* du = (double*)(b + 2 * plane);
    double* dv = (double*)(b + 3 * plane);
    double* dw = (double*)(b + 4 * plane);
    unsigned char* dm = (unsigned char*)(b + 5 * plane);
    cudaStream_t s = ctx->stream;
    const size_t rowb = (size_t)width * sizeof(double);
    OF_CUDA(ctx, cudaMemcpy2DAsync(d1, pitch * 8, img1, step_bytes, rowb, height, cudaMemcpyHostToDevice, s));
    OF_CUDA(ctx, cudaMemcpy2DAsync(d2, pitch * 8, img2, step_bytes, rowb, height, cudaMemcpyHostToDevice, s));
    st = run_batch(ctx, 1, d1, d2, 0, width, height, pitch * 8, n_levels, du, dv, 0, pitch * 8,
                   mask ? dm : nullptr, 0, pitch, nullptr, s);
    if (st) return st;
    OF_CUDA(ctx, cudaMemcpy2DAsync(u, uv_step_bytes, du, pitch * 8, rowb, height, cudaMemcpyDeviceToHost, s));
    OF_CUDA(ctx, cudaMemcpy2DAsync(v, uv_step_bytes, dv, pitch * 8, rowb, height, cudaMemcpyDeviceToHost, s));
    if (mask)
        OF_CUDA(ctx, cudaMemcpy2DAsync(mask, mask_step_bytes, dm, pitch, (size_t)width, height, cudaMemcpyDeviceToHost, s));
    if (ctx->params.write_back_warped && n_levels > 1) {
        // Quirk Q1: the reference leaves the warped level-0 frames in the caller's buffers.
        Plan p = make_plan(width, height, n_levels, 1, false, false);
        ofb::WarpArgs a{};
        a.cu = (const double*)(ctx->scratch.base + p.flow_u_off[1]);
        a.cv = (const double*)(ctx->scratch.base + p.flow_v_off[1]);
        a.c_pitch = p.g[1].pitch;
        a.wc = p.g[1].w;
        a.hc = p.g[1].h;
        a.w = width;
        a.h = height;
        a.src_pitch = pitch;
        a.dst = dw;
        a.dst_pitch = pitch;
        for (int k = 0; k < 2; ++k) {
            a.src = k == 0 ? d1 : d2;
            a.sign = k == 0 ? 1.0f : -1.0f;
            OF_CUDA(ctx, ofb::launch_warp_level(a, s, &ctx->launches));
            OF_CUDA(ctx, cudaMemcpy2DAsync(k == 0 ? img1 : img2, step_bytes, dw, pitch * 8, rowb, height,
                                           cudaMemcpyDeviceToHost, s));
        }
    }
    OF_CUDA(ctx, cudaStreamSynchronize(s));
    return OF_OK;
}

of_status of_lk_single_f64(of_ctx* ctx, const double* img1, const double* img2, int width, int height,
                           size_t step_bytes, double* u, double* v, size_t uv_step_bytes,
                           unsigned char* mask, size_t mask_step_bytes)
{
    // n_levels == 1 never writes the inputs (no warp happens), so the const_cast is safe.
    return of_ctf_f64(ctx, const_cast<double*>(img1), const_cast<double*>(img2), width, height,
                      step_bytes, 1, u, v, uv_step_bytes, mask, mask_step_bytes);
}

}  // extern "C"

// Frame stream, host pointers.  esz = 8: frames are doubles (of_ctf_f64_stream); esz = 1: frames
// are 8-bit and become doubles on the device, v * alpha + beta (of_ctf_u8_stream).
static of_status stream_impl(of_ctx* ctx, const void* frames, size_t esz, double alpha, double beta, int n_frames,
                             size_t frame_stride_bytes, int width, int height, size_t step_bytes, int n_levels,
                             double* u, double* v, size_t uv_pair_stride_bytes, size_t uv_step_bytes,
                             of_gesture_stats* stats)
{
    of_status st = esz == 8 ? check_common(ctx, (const double*)frames, (const double*)frames, width, height, step_bytes, n_levels)
                            : check_common(ctx, (const double*)frames, (const double*)frames, width, height, 8 * step_bytes, n_levels);
    if (st) return st;
    if ((st = reject_write_back(ctx, n_levels, "the stream entry point"))) return st;
    if (n_frames < 2) return fail(ctx, OF_E_BADARG, "n_frames must be >= 2");
    if ((u == nullptr) != (v == nullptr)) return fail(ctx, OF_E_BADARG, "u and v must both be given or both be NULL");
    if (!u && !stats) return fail(ctx, OF_E_BADARG, "nothing to compute: u, v and stats are all NULL");
    if (frame_stride_bytes % esz || frame_stride_bytes < step_bytes * (size_t)(height - 1) + (size_t)width * esz)
        return fail(ctx, OF_E_BADARG, "bad frame_stride_bytes");
    if (u && (uv_step_bytes % 8 || uv_step_bytes < (size_t)width * 8 || uv_pair_stride_bytes % 8))
        return fail(ctx, OF_E_BADARG, "bad uv step/stride");
    OF_CUDA(ctx, cudaSetDevice(ctx->device));

    const int n_pairs = n_frames - 1;
    const size_t pitch = align_up((size_t)width, 16);
    const size_t plane = align_up(pitch * height * sizeof(double), 256);
    // chunk: about 256 MB of frames in flight per buffer, at most 16 pairs -- and small enough
    // that a call is cut into >= 8 chunks when it has the pairs for it: the upload of chunk
    // i+1, the kernels of chunk i and the download of chunk i-1 run concurrently (three
    // streams, two buffers), so only the first upload and the last download are exposed
    int C = (int)(((size_t)256 << 20) / plane);
    if (C > 16) C = 16;
    if (C > (n_pairs + 7) / 8) C = (n_pairs + 7) / 8;
    if (C < 2 && plane <= ((size_t)64 << 20)) C = 2;
    // 8-bit frames in, statistics out: the copies are small, so the chunk is sized for the kernels
    // (full waves of the level kernel) instead of for copy/compute overlap
    if (esz == 1 && !u) C = (int)(((size_t)256 << 20) / plane) < 16 ? (int)(((size_t)256 << 20) / plane) : 16;
    if (const char* e = getenv("OF_B200_STREAM_CHUNK")) {      // development aid (tools/e2e_probe.py)
        const int c = atoi(e);
        if (c >= 1) C = c;
    }
    if (C < 1) C = 1;
    if (C > n_pairs) C = n_pairs;
    const size_t in_bytes = plane * (C + 1);
    const size_t out_bytes = u ? plane * C * 2 : 0;
    const size_t st_bytes = align_up(sizeof(of_gesture_stats) * C, 256);
    const size_t plane8 = align_up((size_t)width * height, 256);       // 8-bit staging frames, dense rows
    const size_t raw_bytes = esz == 1 ? plane8 * (C + 1) : 0;
    const size_t buf = in_bytes + out_bytes + st_bytes + raw_bytes;
    st = arena_reserve(ctx, ctx->staging, 2 * buf);
    if (st) return st;
    if (stats && ctx->h_stats_cap < (size_t)n_pairs) {
        if (ctx->h_stats) cudaFreeHost(ctx->h_stats);
        ctx->h_stats = nullptr;
        ctx->h_stats_cap = 0;
        OF_CUDA(ctx, cudaHostAlloc((void**)&ctx->h_stats, sizeof(of_gesture_stats) * (size_t)n_pairs, cudaHostAllocDefault));
        ctx->h_stats_cap = (size_t)n_pairs;
    }
    const size_t rowb = (size_t)width * sizeof(double);
    const bool dense_in = esz == 8 && step_bytes == rowb && pitch * 8 == rowb && frame_stride_bytes == plane;
    const bool dense_out = u && uv_step_bytes == rowb && pitch * 8 == rowb;

    const int n_chunks = (n_pairs + C - 1) / C;
    for (int i = 0; i < n_chunks; ++i) {
        const int bsel = i & 1;
        char* base = ctx->staging.base + bsel * buf;
        double* d_in = (double*)base;
        double* d_u = (double*)(base + in_bytes);
        double* d_v = (double*)(base + in_bytes + plane * C);
        of_gesture_stats* d_st = (of_gesture_stats*)(base + in_bytes + out_bytes);
        const int p0 = i * C;
        const int np = (p0 + C <= n_pairs) ? C : n_pairs - p0;
        // upload frames p0 .. p0+np; frame p0 of a later chunk is already on the device (the
        // last frame of the previous chunk, in the other buffer): device-to-device copy
        if (i >= 2) OF_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_done[bsel], 0));
        const char* src = (const char*)frames + (size_t)p0 * frame_stride_bytes;
        const int k0 = i > 0 ? 1 : 0;
        if (i > 0) {
            const char* prev_last = ctx->staging.base + (1 - bsel) * buf + plane * (size_t)C;
            OF_CUDA(ctx, cudaMemcpyAsync(d_in, prev_last, plane, cudaMemcpyDeviceToDevice, ctx->copy_stream));
        }
        if (esz == 1) {
            // 8-bit frames: upload (1/8 of the bytes), then convertTo(CV_64FC1, alpha, beta) on the device
            unsigned char* d_raw = (unsigned char*)(base + in_bytes + out_bytes + st_bytes);
            if (step_bytes == (size_t)width && frame_stride_bytes == plane8) {
                // the caller's last frame ends after width*height bytes, not after a padded plane
                OF_CUDA(ctx, cudaMemcpyAsync(d_raw + k0 * plane8, src + (size_t)k0 * frame_stride_bytes,
                                             plane8 * (np - k0) + (size_t)width * height, cudaMemcpyHostToDevice,
                                             ctx->copy_stream));
            } else {
                for (int k = k0; k <= np; ++k)
                    OF_CUDA(ctx, cudaMemcpy2DAsync(d_raw + k * plane8, width, src + (size_t)k * frame_stride_bytes,
                                                   step_bytes, width, height, cudaMemcpyHostToDevice, ctx->copy_stream));
            }
            OF_CUDA(ctx, ofb::launch_u8_to_f64(d_raw + k0 * plane8, plane8, (size_t)width, (double*)((char*)d_in + k0 * plane),
                                               plane / 8, pitch, width, height, np + 1 - k0, alpha, beta,
                                               ctx->copy_stream, &ctx->launches));
        } else if (dense_in) {
            OF_CUDA(ctx, cudaMemcpyAsync((char*)d_in + k0 * plane, src + (size_t)k0 * frame_stride_bytes,
                                         plane * (np - k0) + rowb * height, cudaMemcpyHostToDevice, ctx->copy_stream));
        } else {
            for (int k = k0; k <= np; ++k)
                OF_CUDA(ctx, cudaMemcpy2DAsync((char*)d_in + k * plane, pitch * 8, src + (size_t)k * frame_stride_bytes,
                                               step_bytes, rowb, height, cudaMemcpyHostToDevice, ctx->copy_stream));
        }
        OF_CUDA(ctx, cudaEventRecord(ctx->ev_up[bsel], ctx->copy_stream));
        // compute
        OF_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_up[bsel], 0));
        if (i >= 2) OF_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_down[bsel], 0));
        st = run_batch(ctx, np, d_in, (const double*)((char*)d_in + plane), plane, width, height, pitch * 8,
                       n_levels, u ? d_u : nullptr, u ? d_v : nullptr, plane, pitch * 8, nullptr, 0, 0,
                       stats ? d_st : nullptr, ctx->stream);
        if (st) return st;
        OF_CUDA(ctx, cudaEventRecord(ctx->ev_done[bsel], ctx->stream));
        // download
        OF_CUDA(ctx, cudaStreamWaitEvent(ctx->down_stream, ctx->ev_done[bsel], 0));
        if (u) {
            char* hu = (char*)u + (size_t)p0 * uv_pair_stride_bytes;
            char* hv = (char*)v + (size_t)p0 * uv_pair_stride_bytes;
            if (dense_out && uv_pair_stride_bytes == plane) {
                const size_t nb = plane * (np - 1) + rowb * height;      // the last plane ends at its last pixel
                OF_CUDA(ctx, cudaMemcpyAsync(hu, d_u, nb, cudaMemcpyDeviceToHost, ctx->down_stream));
                OF_CUDA(ctx, cudaMemcpyAsync(hv, d_v, nb, cudaMemcpyDeviceToHost, ctx->down_stream));
            } else {
                for (int k = 0; k < np; ++k) {
                    OF_CUDA(ctx, cudaMemcpy2DAsync(hu + (size_t)k * uv_pair_stride_bytes, uv_step_bytes,
                                                   (char*)d_u + k * plane, pitch * 8, rowb, height,
                                                   cudaMemcpyDeviceToHost, ctx->down_stream));
                    OF_CUDA(ctx, cudaMemcpy2DAsync(hv + (size_t)k * uv_pair_stride_bytes, uv_step_bytes,
                                                   (char*)d_v + k * plane, pitch * 8, rowb, height,
                                                   cudaMemcpyDeviceToHost, ctx->down_stream));
                }
            }
        }
        if (stats)
            OF_CUDA(ctx, cudaMemcpyAsync(ctx->h_stats + p0, d_st, sizeof(of_gesture_stats) * np, cudaMemcpyDeviceToHost,
                                         ctx->down_stream));
        OF_CUDA(ctx, cudaEventRecord(ctx->ev_down[bsel], ctx->down_stream));
    }
    OF_CUDA(ctx, cudaStreamSynchronize(ctx->copy_stream));
    OF_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    OF_CUDA(ctx, cudaStreamSynchronize(ctx->down_stream));
    if (stats) memcpy(stats, ctx->h_stats, sizeof(of_gesture_stats) * (size_t)n_pairs);
    return OF_OK;
}

extern "C" {

of_status of_ctf_f64_stream(of_ctx* ctx, const double* frames, int n_frames, size_t frame_stride_bytes,
                            int width, int height, size_t step_bytes, int n_levels, double* u,
                            double* v, size_t uv_pair_stride_bytes, size_t uv_step_bytes,
                            of_gesture_stats* stats)
{
    return stream_impl(ctx, frames, 8, 1.0, 0.0, n_frames, frame_stride_bytes, width, height, step_bytes, n_levels, u, v,
                       uv_pair_stride_bytes, uv_step_bytes, stats);
}

of_status of_ctf_u8_stream(of_ctx* ctx, const unsigned char* frames, int n_frames, size_t frame_stride_bytes,
                           int width, int height, size_t step_bytes, double alpha, double beta, int n_levels,
                           double* u, double* v, size_t uv_pair_stride_bytes, size_t uv_step_bytes,
                           of_gesture_stats* stats)
{
    if (ctx && (!(alpha == alpha) || !(beta == beta))) return fail(ctx, OF_E_BADARG, "alpha / beta is NaN");
    return stream_impl(ctx, frames, 1, alpha, beta, n_frames, frame_stride_bytes, width, height, step_bytes, n_levels, u, v,
                       uv_pair_stride_bytes, uv_step_bytes, stats);
}

}  // extern "C"

/* ---- allocation-free variant: scratch in a static device array -------------------------------------
 * No counterpart in the reference; added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * of_ctf_f64_dev_static is of_ctf_f64_dev for one pair with the pyramid levels >= 1 of both frames and
 * the accumulated flows of the levels >= 1 (what run_batch keeps in the context's cudaMalloc'ed arena)
 * in the file-scope array below.  Same kernels, same launch sequence, same results.
 *
 * The array belongs to the library, not to a context: every GPU that loads the library holds its
 * OF_STATIC_SCRATCH_BYTES until the process exits, every call on a device shares it, and it keeps the
 * last call's contents.  Nothing in it has to start at zero (every plane is written in full by the call
 * before the call reads it), so nothing is cleared.  Calls that overlap on one GPU (different streams,
 * different contexts, different threads) race on it; serialise them.
 */
static __device__ __align__(256) unsigned char g_static_scratch[OF_STATIC_SCRATCH_BYTES];

namespace {

// In host code `g_static_scratch` is a host-side stand-in for the symbol, not a device address: the
// address comes from cudaGetSymbolAddress, once per device, with that device current.
of_status static_scratch_address(of_ctx* ctx, char** out)
{
    static std::atomic<char*> cache[64];
    int dev = 0;
    OF_CUDA(ctx, cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64) {
        if (char* p = cache[dev].load(std::memory_order_acquire)) {
            *out = p;
            return OF_OK;
        }
    }
    void* p = nullptr;
    OF_CUDA(ctx, cudaGetSymbolAddress(&p, g_static_scratch));
    if (!p) return fail(ctx, OF_E_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
    if (dev >= 0 && dev < 64) cache[dev].store((char*)p, std::memory_order_release);
    *out = (char*)p;
    return OF_OK;
}

}  // namespace

extern "C" {

of_status of_ctf_f64_dev_static(of_ctx* ctx, const double* img1, const double* img2, int width, int height,
                                size_t step_bytes, int n_levels, double* u, double* v, size_t uv_step_bytes,
                                void* stream)
{
    of_status st = check_common(ctx, img1, img2, width, height, step_bytes, n_levels);
    if (st) return st;
    if (!u || !v) return fail(ctx, OF_E_BADARG, "null flow pointer");
    if (uv_step_bytes % sizeof(double) || uv_step_bytes < (size_t)width * sizeof(double))
        return fail(ctx, OF_E_BADARG, "uv_step_bytes %zu invalid for width %d", uv_step_bytes, width);
    if ((st = reject_write_back(ctx, n_levels, "of_ctf_f64_dev_static"))) return st;
    const Plan p = make_plan(width, height, n_levels, 1, false, false);
    if (p.total > (size_t)OF_STATIC_SCRATCH_BYTES)
        return fail(ctx, OF_E_BADARG,
                    "of_ctf_f64_dev_static: %dx%d with %d levels needs %zu bytes of scratch, the static array holds %zu "
                    "(OF_STATIC_SCRATCH_BYTES); use of_ctf_f64_dev",
                    width, height, n_levels, p.total, (size_t)OF_STATIC_SCRATCH_BYTES);
    OF_CUDA(ctx, cudaSetDevice(ctx->device));
    char* base = nullptr;
    if ((st = static_scratch_address(ctx, &base))) return st;
    cudaStream_t s = pick(ctx, stream);
    const size_t in_pitch = step_bytes / sizeof(double);
    const int L = n_levels;

    // Gaussian pyramids (getGaussianPyramid :566-575): frame 1 in image slot 0, frame 2 in slot 1 of every level
    for (int l = 1; l < L; ++l) {
        const LevelGeom& gs = p.g[l - 1];
        const LevelGeom& gd = p.g[l];
        ofb::PyrDownArgs a{};
        a.w = gs.w;
        a.h = gs.h;
        a.dst = (double*)(base + p.pyr_off[l]);
        a.dst_pitch = gd.pitch;
        a.dst_img_stride = gd.img_stride;
        if (l == 1) {
            a.src_pitch = in_pitch;
            a.n_images = 1;
            a.src = img1;
            OF_CUDA(ctx, ofb::launch_pyr_down(a, s, &ctx->launches));
            a.src = img2;
            a.dst += gd.img_stride;
            OF_CUDA(ctx, ofb::launch_pyr_down(a, s, &ctx->launches));
        } else {
            a.src = (const double*)(base + p.pyr_off[l - 1]);
            a.src_pitch = gs.pitch;
            a.src_img_stride = gs.img_stride;
            a.n_images = 2;
            OF_CUDA(ctx, ofb::launch_pyr_down(a, s, &ctx->launches));
        }
    }
    // coarse to fine (:610-703), one fused launch per level
    for (int l = L - 1; l >= 0; --l) {
        const LevelGeom& g = p.g[l];
        ofb::LevelArgs a{};
        a.w = g.w;
        a.h = g.h;
        a.radius = ctx->params.win_radius;
        a.det_eps = ctx->params.det_eps;
        a.mag_thresh = ctx->params.mag_thresh;
        a.mag2_thresh = mag2_threshold(ctx->params.mag_thresh);
        if (l == 0) {
            a.A = img1;
            a.B = img2;
            a.img_pitch = in_pitch;
        } else {
            const double* pyr = (const double*)(base + p.pyr_off[l]);
            a.A = pyr;
            a.B = pyr + g.img_stride;
            a.img_pitch = g.pitch;
            a.a_pair_stride = a.b_pair_stride = g.img_stride;
        }
        if (l < L - 1) {
            const LevelGeom& gc = p.g[l + 1];
            a.cu = (const double*)(base + p.flow_u_off[l + 1]);
            a.cv = (const double*)(base + p.flow_v_off[l + 1]);
            a.c_pitch = gc.pitch;
            a.c_pair_stride = gc.img_stride;
            a.wc = gc.w;
            a.hc = gc.h;
        }
        if (l == 0) {
            a.u = u;
            a.v = v;
            a.uv_pitch = uv_step_bytes / sizeof(double);
        } else {
            a.u = (double*)(base + p.flow_u_off[l]);
            a.v = (double*)(base + p.flow_v_off[l]);
            a.uv_pitch = g.pitch;
            a.uv_pair_stride = g.img_stride;
        }
        OF_CUDA(ctx, launch_lk_level(ctx, a, 1, s));
    }
    return OF_OK;
}

}  // extern "C"
