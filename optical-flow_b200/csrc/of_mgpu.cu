// Single-process multi-GPU host path (include/of_b200.h, of_mgpu_*): frame pairs of a stream shard
// across the GPUs of one box.  One of_ctx per device, one host thread per device
// inside each call, contiguous blocks of pairs (a device uploads every interior frame of its block
// once and reuses its pyramid for both pairs that share it), flow fields and gesture statistics
// written straight into the caller's arrays.  A frame is never split and the devices exchange
// nothing: there is no collective on this path.
#include <cuda_runtime.h>

#include <cstdio>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "../../include/of_b200.h"

struct of_mgpu {
    std::vector<int> devices;
    std::vector<of_ctx*> ctx;
    std::string err;
};

namespace {

std::string g_mgpu_create_error;

void partition(int n_pairs, int n_dev, int idx, int* first, int* last)
{
    const int base = n_pairs / n_dev, rem = n_pairs % n_dev;
    *first = idx * base + (idx < rem ? idx : rem);
    *last = *first + base + (idx < rem ? 1 : 0);
}

// esz = 8: double frames (of_ctf_f64_stream); esz = 1: 8-bit frames (of_ctf_u8_stream)
of_status mgpu_stream(of_mgpu* m, const void* frames, size_t esz, double alpha, double beta, int n_frames,
                      size_t frame_stride_bytes, int width, int height, size_t step_bytes, int n_levels, double* u,
                      double* v, size_t uv_pair_stride_bytes, size_t uv_step_bytes, of_gesture_stats* stats)
{
    if (!m) return OF_E_BADARG;
    if (!frames || n_frames < 2) {
        m->err = "frames is NULL or n_frames < 2";
        return OF_E_BADARG;
    }
    const int n_pairs = n_frames - 1;
    const int n_dev = (int)m->ctx.size();
    std::vector<of_status> st(n_dev, OF_OK);
    std::vector<std::thread> th;
    th.reserve(n_dev);
    for (int d = 0; d < n_dev; ++d) {
        int p0, p1;
        partition(n_pairs, n_dev, d, &p0, &p1);
        if (p1 <= p0) continue;                              // more devices than pairs
        th.emplace_back([=, &st]() {
            const char* f = (const char*)frames + (size_t)p0 * frame_stride_bytes;
            double* du = u ? (double*)((char*)u + (size_t)p0 * uv_pair_stride_bytes) : nullptr;
            double* dv = v ? (double*)((char*)v + (size_t)p0 * uv_pair_stride_bytes) : nullptr;
            of_gesture_stats* ds = stats ? stats + p0 : nullptr;
            const int nf = p1 - p0 + 1;
            st[d] = esz == 8 ? of_ctf_f64_stream(m->ctx[d], (const double*)f, nf, frame_stride_bytes, width, height,
                                                 step_bytes, n_levels, du, dv, uv_pair_stride_bytes, uv_step_bytes, ds)
                             : of_ctf_u8_stream(m->ctx[d], (const unsigned char*)f, nf, frame_stride_bytes, width, height,
                                                step_bytes, alpha, beta, n_levels, du, dv, uv_pair_stride_bytes,
                                                uv_step_bytes, ds);
        });
    }
    for (std::thread& t : th) t.join();
    for (int d = 0; d < n_dev; ++d) {
        if (st[d] != OF_OK) {
            char buf[64];
            snprintf(buf, sizeof buf, "device %d: ", m->devices[d]);
            m->err = std::string(buf) + of_last_error_string(m->ctx[d]);
            return st[d];
        }
    }
    return OF_OK;
}

}  // namespace

extern "C" {

of_status of_mgpu_create(const int* devices, int n_devices, const of_params* params, of_mgpu** out)
{
    if (!out) return OF_E_BADARG;
    *out = nullptr;
    int visible = 0;
    if (cudaGetDeviceCount(&visible) != cudaSuccess || visible == 0) {
        cudaGetLastError();
        g_mgpu_create_error = "no usable CUDA device; this library has no CPU fallback";
        return OF_E_CUDA;
    }
    if (n_devices < 0 || (n_devices > 0 && !devices)) {
        g_mgpu_create_error = "bad device list";
        return OF_E_BADARG;
    }
    of_mgpu* m = new (std::nothrow) of_mgpu();
    if (!m) return OF_E_NOMEM;
    if (n_devices == 0) {
        for (int d = 0; d < visible; ++d) m->devices.push_back(d);
    } else {
        for (int i = 0; i < n_devices; ++i) {
            for (int j = 0; j < i; ++j)
                if (devices[j] == devices[i]) {
                    g_mgpu_create_error = "a device is listed twice";
                    delete m;
                    return OF_E_BADARG;
                }
            m->devices.push_back(devices[i]);
        }
    }
    for (int d : m->devices) {
        of_ctx* c = nullptr;
        const of_status st = of_ctx_create(d, params, &c);
        if (st != OF_OK) {
            g_mgpu_create_error = of_last_error_string(nullptr);
            of_mgpu_destroy(m);
            return st;
        }
        m->ctx.push_back(c);
    }
    *out = m;
    return OF_OK;
}

void of_mgpu_destroy(of_mgpu* m)
{
    if (!m) return;
    for (of_ctx* c : m->ctx) of_ctx_destroy(c);
    delete m;
}

int of_mgpu_device_count(const of_mgpu* m) { return m ? (int)m->ctx.size() : 0; }

of_ctx* of_mgpu_context(of_mgpu* m, int index) { return m && index >= 0 && index < (int)m->ctx.size() ? m->ctx[index] : nullptr; }

const char* of_mgpu_last_error_string(const of_mgpu* m) { return m ? m->err.c_str() : g_mgpu_create_error.c_str(); }

of_status of_mgpu_partition(int n_pairs, int n_devices, int index, int* first, int* last)
{
    if (n_pairs < 0 || n_devices < 1 || index < 0 || index >= n_devices || !first || !last) return OF_E_BADARG;
    partition(n_pairs, n_devices, index, first, last);
    return OF_OK;
}

of_status of_mgpu_ctf_f64_stream(of_mgpu* m, const double* frames, int n_frames, size_t frame_stride_bytes, int width,
                                 int height, size_t step_bytes, int n_levels, double* u, double* v,
                                 size_t uv_pair_stride_bytes, size_t uv_step_bytes, of_gesture_stats* stats)
{
    return mgpu_stream(m, frames, 8, 1.0, 0.0, n_frames, frame_stride_bytes, width, height, step_bytes, n_levels, u, v,
                       uv_pair_stride_bytes, uv_step_bytes, stats);
}

of_status of_mgpu_ctf_u8_stream(of_mgpu* m, const unsigned char* frames, int n_frames, size_t frame_stride_bytes,
                                int width, int height, size_t step_bytes, double alpha, double beta, int n_levels,
                                double* u, double* v, size_t uv_pair_stride_bytes, size_t uv_step_bytes,
                                of_gesture_stats* stats)
{
    return mgpu_stream(m, frames, 1, alpha, beta, n_frames, frame_stride_bytes, width, height, step_bytes, n_levels, u, v,
                       uv_pair_stride_bytes, uv_step_bytes, stats);
}

}  // extern "C"
