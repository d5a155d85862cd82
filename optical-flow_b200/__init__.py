"""optical-flow_b200 -- B200-native dense pyramidal Lucas-Kanade optical flow.

Python host side of the C-ABI library ``libof_b200.so`` (``include/of_b200.h``).  It
mirrors the reference's three entry points by name
(``/root/reference/ParallelVideoPyr.cpp``: ``getMaxLayer`` :545, ``getLucasKanadeOpticalFlow``
:327, ``coarseToFineEstimation`` :578) over numpy arrays (host-pointer calls) and torch CUDA
tensors (device-pointer calls).  Python is plumbing only: every flow value is produced by the
hand-written sm_100a kernels behind the C ABI.  There is NO CPU fallback -- if the shared
library is missing or no CUDA device is usable, calls raise.

The directory name contains a hyphen, so import it through ``of_b200_loader`` at the repo
root (``from of_b200_loader import ofb``).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("OF_B200_LIB") or os.path.join(_HERE, "libof_b200.so")   # env: A/B builds (development)

OF_OK, OF_E_BADARG, OF_E_NOT_DIVISIBLE, OF_E_CUDA, OF_E_NOMEM = 0, 1, 2, 3, 4

EXPORTS = (
    "of_params_default", "of_get_max_layer", "of_ctx_create", "of_ctx_destroy",
    "of_ctx_set_params", "of_last_error_string", "of_abi_version", "of_lk_single_f64",
    "of_ctf_f64", "of_ctf_f64_stream", "of_lk_single_f64_dev", "of_ctf_f64_dev",
    "of_ctf_f64_batch_dev", "of_gesture_stats_dev", "of_pyr_down_f64_dev",
    "of_ctx_synchronize", "of_ctx_launch_count", "of_ctx_set_timing", "of_ctx_get_timing",
    "of_bgr2gray_u8_dev", "of_live_diff_f64_dev", "of_live_diff_f64", "of_ctf_u8_stream",
    "of_ctx_get_level_f64", "of_ctx_graph_replay_count", "of_ctf_f64_dev_static",
    "of_resize_cubic_u8_dev", "of_gaussian_blur_u8_dev", "of_live_params_default", "of_live_step_u8", "of_live_reset",
    "of_mgpu_create", "of_mgpu_destroy", "of_mgpu_device_count", "of_mgpu_context", "of_mgpu_last_error_string",
    "of_mgpu_partition", "of_mgpu_ctf_f64_stream", "of_mgpu_ctf_u8_stream",
)


class OfParams(C.Structure):
    _fields_ = [("win_radius", C.c_int), ("det_eps", C.c_double), ("write_back_warped", C.c_int),
                ("mag_thresh", C.c_double), ("count_thresh", C.c_int)]


class OfLiveParams(C.Structure):
    _fields_ = [("work_width", C.c_int), ("work_height", C.c_int), ("blur_ksize", C.c_int), ("blur_sigma", C.c_double),
                ("coeff_bits", C.c_int), ("learning_rate", C.c_double), ("diff_thresh", C.c_int),
                ("morph_iterations", C.c_int), ("n_levels", C.c_int)]


class OfGestureStats(C.Structure):
    _fields_ = [("sum_x", C.c_double), ("sum_y", C.c_double), ("count", C.c_double),
                ("detected", C.c_double)]


STATS_DTYPE = np.dtype([("sum_x", "<f8"), ("sum_y", "<f8"), ("count", "<f8"), ("detected", "<f8")])


class OfError(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__("of_b200 error %d: %s" % (status, msg))
        self.status = status


class NotDivisibleError(OfError, ValueError):
    """W or H not divisible by 2^(nLevels-1): the reference throws cv::Exception (Q2)."""


class BadArgError(OfError, ValueError):
    pass


_lib = None


def lib():
    """Loads libof_b200.so; raises (never falls back) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "%s is missing: build it with `python optical-flow_b200/build.py` "
            "(__graft_entry__.build()).  There is no CPU fallback." % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, sz, ci, dbl = C.c_void_p, C.c_size_t, C.c_int, C.c_double
    L.of_params_default.argtypes = [C.POINTER(OfParams)]
    L.of_params_default.restype = None
    L.of_get_max_layer.argtypes = [ci, ci]
    L.of_ctx_create.argtypes = [ci, C.POINTER(OfParams), C.POINTER(vp)]
    L.of_ctx_destroy.argtypes = [vp]
    L.of_ctx_destroy.restype = None
    L.of_ctx_set_params.argtypes = [vp, C.POINTER(OfParams)]
    L.of_last_error_string.argtypes = [vp]
    L.of_last_error_string.restype = C.c_char_p
    L.of_abi_version.argtypes = []
    L.of_lk_single_f64.argtypes = [vp, vp, vp, ci, ci, sz, vp, vp, sz, vp, sz]
    L.of_ctf_f64.argtypes = [vp, vp, vp, ci, ci, sz, ci, vp, vp, sz, vp, sz]
    L.of_ctf_f64_stream.argtypes = [vp, vp, ci, sz, ci, ci, sz, ci, vp, vp, sz, sz, vp]
    L.of_ctf_u8_stream.argtypes = [vp, vp, ci, sz, ci, ci, sz, dbl, dbl, ci, vp, vp, sz, sz, vp]
    L.of_lk_single_f64_dev.argtypes = [vp, vp, vp, ci, ci, sz, vp, vp, sz, vp, sz, vp]
    L.of_ctf_f64_dev.argtypes = [vp, vp, vp, ci, ci, sz, ci, vp, vp, sz, vp, sz, vp]
    L.of_ctf_f64_dev_static.argtypes = [vp, vp, vp, ci, ci, sz, ci, vp, vp, sz, vp]
    L.of_ctf_f64_batch_dev.argtypes = [vp, ci, vp, vp, sz, ci, ci, sz, ci, vp, vp, sz, sz,
                                       vp, sz, sz, vp, vp]
    L.of_gesture_stats_dev.argtypes = [vp, ci, vp, vp, sz, ci, ci, sz, vp, vp]
    L.of_pyr_down_f64_dev.argtypes = [vp, vp, ci, ci, sz, vp, sz, vp]
    L.of_ctx_get_level_f64.argtypes = [vp, ci, ci, ci, vp, sz, C.POINTER(ci), C.POINTER(ci)]
    L.of_bgr2gray_u8_dev.argtypes = [vp, vp, ci, ci, sz, vp, sz, ci, vp]
    L.of_live_diff_f64_dev.argtypes = [vp, vp, vp, ci, ci, sz, dbl, ci, ci, vp, sz, vp]
    L.of_live_diff_f64.argtypes = [vp, vp, vp, ci, ci, ci, sz, ci, dbl, ci, ci, vp, sz]
    L.of_mgpu_create.argtypes = [C.POINTER(ci), ci, C.POINTER(OfParams), C.POINTER(vp)]
    L.of_mgpu_destroy.argtypes = [vp]
    L.of_mgpu_destroy.restype = None
    L.of_mgpu_device_count.argtypes = [vp]
    L.of_mgpu_context.argtypes = [vp, ci]
    L.of_mgpu_context.restype = vp
    L.of_mgpu_last_error_string.argtypes = [vp]
    L.of_mgpu_last_error_string.restype = C.c_char_p
    L.of_mgpu_partition.argtypes = [ci, ci, ci, C.POINTER(ci), C.POINTER(ci)]
    L.of_mgpu_ctf_f64_stream.argtypes = [vp, vp, ci, sz, ci, ci, sz, ci, vp, vp, sz, sz, vp]
    L.of_mgpu_ctf_u8_stream.argtypes = [vp, vp, ci, sz, ci, ci, sz, dbl, dbl, ci, vp, vp, sz, sz, vp]
    L.of_ctx_synchronize.argtypes = [vp]
    L.of_ctx_launch_count.argtypes = [vp]
    L.of_ctx_launch_count.restype = C.c_ulonglong
    L.of_resize_cubic_u8_dev.argtypes = [vp, vp, ci, ci, sz, ci, vp, ci, ci, sz, vp]
    L.of_gaussian_blur_u8_dev.argtypes = [vp, vp, ci, ci, sz, ci, ci, dbl, ci, vp, sz, vp]
    L.of_live_params_default.argtypes = [C.POINTER(OfLiveParams)]
    L.of_live_params_default.restype = None
    L.of_live_step_u8.argtypes = [vp, C.POINTER(OfLiveParams), vp, ci, ci, sz, vp, vp, vp, sz, C.POINTER(ci)]
    L.of_live_reset.argtypes = [vp]
    L.of_ctx_graph_replay_count.argtypes = [vp]
    L.of_ctx_graph_replay_count.restype = C.c_ulonglong
    L.of_ctx_set_timing.argtypes = [vp, ci]
    L.of_ctx_get_timing.argtypes = [vp, ci, vp, vp, vp, vp]
    _ = dbl
    _lib = L
    return L


def default_params() -> OfParams:
    p = OfParams()
    lib().of_params_default(C.byref(p))
    return p


def getMaxLayer(img_or_width, height=None) -> int:
    """``int getMaxLayer(Mat &img)`` (ParallelVideoPyr.cpp:545-564).  Accepts an array
    (rows, cols) like the reference, or (width, height)."""
    if height is None:
        h, w = img_or_width.shape[-2:]
    else:
        w, h = int(img_or_width), int(height)
    return int(lib().of_get_max_layer(w, h))


def _np_f64_2d(a, name):
    a = np.asarray(a)
    if a.dtype != np.float64 or a.ndim != 2:
        raise BadArgError(OF_E_BADARG, "%s must be a 2-D float64 array (CV_64FC1)" % name)
    if a.strides[1] != 8 or a.strides[0] < a.shape[1] * 8 or a.strides[0] % 8:
        a = np.ascontiguousarray(a)
    return a


def _stream_handle(stream, device):
    """cudaStream_t of a torch stream for the C ABI.  torch's default stream has handle 0,
    which the ABI reads as "the context's own stream", so it is passed as cudaStreamLegacy
    (0x1) -- the same stream, named explicitly."""
    import torch
    h = (stream or torch.cuda.current_stream(device)).cuda_stream
    return h if h else 1


class Context:
    """One ``of_ctx`` (one CUDA device).  Methods carry the reference's function names."""

    def __init__(self, device: int = 0, win_radius: int = 1, det_eps: float = 0.0,
                 write_back_warped: bool = False, mag_thresh: float = 20.0, count_thresh: int = 500):
        L = lib()
        p = default_params()
        p.win_radius, p.det_eps = int(win_radius), float(det_eps)
        p.write_back_warped = int(bool(write_back_warped))
        p.mag_thresh, p.count_thresh = float(mag_thresh), int(count_thresh)
        self._params = p
        self._h = C.c_void_p()
        st = L.of_ctx_create(int(device), C.byref(p), C.byref(self._h))
        if st != OF_OK:
            raise OfError(st, (L.of_last_error_string(None) or b"").decode())
        self.device = int(device)

    # -- lifetime ---------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().of_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, st):
        if st == OF_OK:
            return
        msg = (lib().of_last_error_string(self._h) or b"").decode()
        if st == OF_E_NOT_DIVISIBLE:
            raise NotDivisibleError(st, msg)
        if st == OF_E_BADARG:
            raise BadArgError(st, msg)
        raise OfError(st, msg)

    def set_params(self, **kw):
        for k, v in kw.items():
            if not hasattr(self._params, k):
                raise BadArgError(OF_E_BADARG, "unknown parameter %s" % k)
            setattr(self._params, k, v)
        self._check(lib().of_ctx_set_params(self._h, C.byref(self._params)))

    @property
    def launch_count(self) -> int:
        return int(lib().of_ctx_launch_count(self._h))

    @property
    def graph_replays(self) -> int:
        return int(lib().of_ctx_graph_replay_count(self._h))

    def synchronize(self):
        self._check(lib().of_ctx_synchronize(self._h))

    def set_timing(self, enable: bool):
        self._check(lib().of_ctx_set_timing(self._h, int(bool(enable))))

    def get_level(self, what: str, level: int, pair: int = 0):
        """Parity triage: plane ``what`` ('frame1', 'frame2', 'u', 'v') of pyramid level ``level`` >= 1
        of pair ``pair`` as left by the most recent pyramid call (numpy array)."""
        code = {"frame1": 0, "frame2": 1, "u": 2, "v": 3}[what]
        w, h = C.c_int(0), C.c_int(0)
        self._check(lib().of_ctx_get_level_f64(self._h, code, int(level), int(pair), None, 0, C.byref(w), C.byref(h)))
        out = np.empty((h.value, w.value), np.float64)
        self._check(lib().of_ctx_get_level_f64(self._h, code, int(level), int(pair), out.ctypes.data_as(C.c_void_p),
                                               out.strides[0], C.byref(w), C.byref(h)))
        return out

    def get_timing(self):
        """[(kind, level, ms)] of every launch since set_timing(True) (at most 8192);
        kind 0 = pyrDown producing `level`, 1 = fused level kernel."""
        n = C.c_int(0)
        cap = 8192
        kinds = (C.c_int * cap)()
        levels = (C.c_int * cap)()
        ms = (C.c_float * cap)()
        self._check(lib().of_ctx_get_timing(self._h, cap, kinds, levels, ms, C.byref(n)))
        return [(kinds[i], levels[i], ms[i]) for i in range(n.value)]

    # -- host-pointer calls (numpy) -----------------------------------------------------
    def getLucasKanadeOpticalFlow(self, img1, img2, want_mask: bool = False):
        """``getLucasKanadeOpticalFlow(img1, img2, u, v)`` (:327-542).  Returns (u, v[, mask])
        as fresh arrays, like the reference rebinding ``u``/``v`` (Q4)."""
        return self._host(img1, img2, 1, want_mask, single=True)

    def coarseToFineEstimation(self, img1, img2, nLevels: int, want_mask: bool = False):
        """``coarseToFineEstimation(img1, img2, u, v, nLevels)`` (:578-725).  With
        ``write_back_warped`` the numpy inputs are overwritten like the reference (Q1)."""
        return self._host(img1, img2, int(nLevels), want_mask, single=False)

    def _host(self, img1, img2, n_levels, want_mask, single):
        a, b = _np_f64_2d(img1, "img1"), _np_f64_2d(img2, "img2")
        if a.shape != b.shape:
            raise BadArgError(OF_E_BADARG, "frame sizes differ")
        if a.strides[0] != b.strides[0]:
            a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
        h, w = a.shape
        u = np.empty((h, w), np.float64)
        v = np.empty((h, w), np.float64)
        m = np.empty((h, w), np.uint8) if want_mask else None
        L = lib()
        if single:
            st = L.of_lk_single_f64(self._h, a.ctypes.data, b.ctypes.data, w, h, a.strides[0],
                                    u.ctypes.data, v.ctypes.data, w * 8,
                                    m.ctypes.data if want_mask else None, w)
        else:
            wb = bool(self._params.write_back_warped)
            if wb and (a is not img1 or b is not img2 or not a.flags.writeable or not b.flags.writeable):
                raise BadArgError(OF_E_BADARG, "write_back_warped needs writable float64 arrays passed as-is")
            st = L.of_ctf_f64(self._h, a.ctypes.data, b.ctypes.data, w, h, a.strides[0], n_levels,
                              u.ctypes.data, v.ctypes.data, w * 8,
                              m.ctypes.data if want_mask else None, w)
        self._check(st)
        return (u, v, m.astype(bool)) if want_mask else (u, v)

    def flow_stream(self, frames, nLevels: int, want_flow: bool = True, want_stats: bool = True,
                    out_u=None, out_v=None, alpha: float = 1.0 / 255, beta: float = 0.0):
        """Host-pointer frame stream: ``frames`` (n, H, W) float64 (pinned memory makes the
        copies asynchronous) -> flows of the n-1 consecutive pairs and/or their gesture
        statistics (the live loop :769-905 without camera and window)."""
        f = np.asarray(frames)
        if f.dtype not in (np.float64, np.uint8) or f.ndim != 3 or not f.flags.c_contiguous:
            raise BadArgError(OF_E_BADARG, "frames must be a C-contiguous (n, H, W) float64 or uint8 array")
        n, h, w = f.shape
        u = v = None
        if want_flow:
            u = out_u if out_u is not None else np.empty((n - 1, h, w), np.float64)
            v = out_v if out_v is not None else np.empty((n - 1, h, w), np.float64)
        stats = np.zeros(n - 1, STATS_DTYPE) if want_stats else None
        if f.dtype == np.uint8:      # 8-bit frames, converted on the device (convertTo(CV_64FC1, alpha, beta))
            st = lib().of_ctf_u8_stream(self._h, f.ctypes.data, n, h * w, w, h, w, float(alpha), float(beta),
                                        int(nLevels), u.ctypes.data if want_flow else None,
                                        v.ctypes.data if want_flow else None, h * w * 8, w * 8,
                                        stats.ctypes.data if want_stats else None)
            self._check(st)
            return u, v, stats
        st = lib().of_ctf_f64_stream(self._h, f.ctypes.data, n, h * w * 8, w, h, w * 8, int(nLevels),
                                     u.ctypes.data if want_flow else None,
                                     v.ctypes.data if want_flow else None, h * w * 8, w * 8,
                                     stats.ctypes.data if want_stats else None)
        self._check(st)
        return u, v, stats

    # -- device-pointer calls (torch CUDA tensors) ----------------------------------------
    @staticmethod
    def _t2d(t, name):
        if t.dtype.__str__() != "torch.float64" or t.dim() != 2 or not t.is_cuda or t.stride(1) != 1:
            raise BadArgError(OF_E_BADARG, "%s must be a 2-D float64 CUDA tensor with unit column stride" % name)
        return t

    def coarseToFineEstimation_dev(self, img1, img2, nLevels: int, u, v, mask=None, stream=None):
        """Device-resident ``coarseToFineEstimation``: all arguments are torch CUDA tensors
        (u, v pre-allocated, as the C ABI requires); asynchronous on ``stream`` (a
        ``torch.cuda.Stream``; default: torch's current stream)."""
        import torch
        a, b = self._t2d(img1, "img1"), self._t2d(img2, "img2")
        self._t2d(u, "u"), self._t2d(v, "v")
        if a.shape != b.shape or u.shape != a.shape or v.shape != a.shape or a.stride(0) != b.stride(0) \
                or u.stride(0) != v.stride(0):
            raise BadArgError(OF_E_BADARG, "shape/stride mismatch")
        h, w = a.shape
        s = _stream_handle(stream, a.device)
        st = lib().of_ctf_f64_dev(self._h, a.data_ptr(), b.data_ptr(), w, h, a.stride(0) * 8, int(nLevels),
                                  u.data_ptr(), v.data_ptr(), u.stride(0) * 8,
                                  mask.data_ptr() if mask is not None else None,
                                  mask.stride(0) if mask is not None else 0, s)
        self._check(st)

    def getLucasKanadeOpticalFlow_dev(self, img1, img2, u, v, mask=None, stream=None):
        self.coarseToFineEstimation_dev(img1, img2, 1, u, v, mask, stream)

    def flow_batch_dev(self, frames, nLevels: int, u=None, v=None, stats=None, mask=None,
                       stream=None, pairs=None):
        """Device-resident batch.  ``frames``: (n, H, W) float64 CUDA tensor.  Default pairing
        is consecutive frames (k, k+1) -> n-1 pairs (pyramids shared between neighbours);
        ``pairs="split"`` treats frames[:n/2] as first and frames[n/2:] as second frames.
        u, v: (n_pairs, H, W) tensors or None; stats: (n_pairs, 4) float64 tensor or None."""
        import torch
        f = frames
        if f.dtype != torch.float64 or f.dim() != 3 or not f.is_cuda or not f.is_contiguous():
            raise BadArgError(OF_E_BADARG, "frames must be a contiguous (n, H, W) float64 CUDA tensor")
        n, h, w = f.shape
        fb = h * w * 8
        if pairs == "split":
            npairs = n // 2
            p1, p2 = f.data_ptr(), f.data_ptr() + npairs * fb
        else:
            npairs = n - 1
            p1, p2 = f.data_ptr(), f.data_ptr() + fb
        for t, nm in ((u, "u"), (v, "v")):
            if t is not None and (t.dtype != torch.float64 or tuple(t.shape) != (npairs, h, w)
                                  or not t.is_contiguous() or not t.is_cuda):
                raise BadArgError(OF_E_BADARG, "%s must be a contiguous (%d,%d,%d) float64 CUDA tensor" % (nm, npairs, h, w))
        if stats is not None and (stats.dtype != torch.float64 or tuple(stats.shape) != (npairs, 4)
                                  or not stats.is_contiguous()):
            raise BadArgError(OF_E_BADARG, "stats must be a contiguous (n_pairs, 4) float64 CUDA tensor")
        if mask is not None and (mask.dtype != torch.uint8 or tuple(mask.shape) != (npairs, h, w)
                                 or not mask.is_contiguous()):
            raise BadArgError(OF_E_BADARG, "mask must be a contiguous (n_pairs, H, W) uint8 CUDA tensor")
        s = _stream_handle(stream, f.device)
        st = lib().of_ctf_f64_batch_dev(
            self._h, npairs, p1, p2, fb, w, h, w * 8, int(nLevels),
            u.data_ptr() if u is not None else None, v.data_ptr() if v is not None else None, fb, w * 8,
            mask.data_ptr() if mask is not None else None, h * w, w,
            stats.data_ptr() if stats is not None else None, s)
        self._check(st)
        return npairs

    def gesture_stats_dev(self, u, v, stats, stream=None):
        import torch
        npairs, h, w = u.shape
        s = _stream_handle(stream, u.device)
        self._check(lib().of_gesture_stats_dev(self._h, npairs, u.data_ptr(), v.data_ptr(), h * w * 8,
                                               w, h, w * 8, stats.data_ptr(), s))

    def pyr_down_dev(self, src, dst, stream=None):
        import torch
        h, w = src.shape
        s = _stream_handle(stream, src.device)
        self._check(lib().of_pyr_down_f64_dev(self._h, src.data_ptr(), w, h, src.stride(0) * 8,
                                              dst.data_ptr(), dst.stride(0) * 8, s))

    # ---- live-loop pre-processing (ParallelVideoPyr.cpp:791-813) ----
    def live_diff(self, cur, prev, learning_rate: float = 0.3, diff_thresh: int = 10,
                  morph_iterations: int = 2, coeff_bits: int = 14):
        """The LK input of one live-loop iteration from two consecutive 8-bit frames (host arrays,
        (H, W) gray or (H, W, 3) BGR): cvtColor :791, ``cur - 0.3 * prev`` :800, threshold :803,
        Sobel x + Sobel y :808-811, dilate x2 / erode x2 :812-813.  Returns an (H, W) float64 array."""
        cur = np.ascontiguousarray(cur)
        prev = np.ascontiguousarray(prev)
        if cur.dtype != np.uint8 or prev.dtype != np.uint8 or cur.shape != prev.shape or cur.ndim not in (2, 3):
            raise BadArgError(OF_E_BADARG, "cur / prev must be uint8 arrays of the same (H, W) or (H, W, 3) shape")
        ch = 1 if cur.ndim == 2 else cur.shape[2]
        h, w = cur.shape[:2]
        out = np.empty((h, w), np.float64)
        self._check(lib().of_live_diff_f64(self._h, cur.ctypes.data, prev.ctypes.data, ch, w, h, w * ch,
                                           coeff_bits, learning_rate, diff_thresh, morph_iterations,
                                           out.ctypes.data, w * 8))
        return out

    def resize_cubic_dev(self, src, dst, stream=None):
        """cv::resize(INTER_CUBIC) :782 on device tensors: src (H, W[, 3]) uint8 -> dst (H2, W2[, 3]) uint8."""
        cn = 1 if src.dim() == 2 else src.shape[2]
        s = _stream_handle(stream, src.device)
        self._check(lib().of_resize_cubic_u8_dev(self._h, src.data_ptr(), src.shape[1], src.shape[0], src.stride(0), cn,
                                                 dst.data_ptr(), dst.shape[1], dst.shape[0], dst.stride(0), s))

    def gaussian_blur_dev(self, src, dst, ksize: int = 9, sigma: float = 1.5, gray_coeff_bits: int = 0, stream=None):
        """cv::GaussianBlur :790 on device tensors (uint8, (H, W) or (H, W, 3)); gray_coeff_bits = 14 / 15 fuses
        cvtColor(BGR2GRAY) :791 and writes an (H, W) gray image."""
        cn = 1 if src.dim() == 2 else src.shape[2]
        s = _stream_handle(stream, src.device)
        self._check(lib().of_gaussian_blur_u8_dev(self._h, src.data_ptr(), src.shape[1], src.shape[0], src.stride(0), cn,
                                                  int(ksize), float(sigma), int(gray_coeff_bits), dst.data_ptr(),
                                                  dst.stride(0), s))

    def live_reset(self):
        self._check(lib().of_live_reset(self._h))

    def live_step(self, frame_bgr, want_flow: bool = False, **kw):
        """One iteration of the reference's live loop (main(), :769-905) on a captured 8-bit BGR frame (host
        array (H, W, 3)); keyword arguments override of_live_params (work_width, work_height, blur_ksize,
        blur_sigma, coeff_bits, learning_rate, diff_thresh, morph_iterations, n_levels).
        Returns (stage, stats, u, v): stage 0 / 1 = warm-up iterations (no result), 2 = stats valid."""
        f = np.ascontiguousarray(frame_bgr)
        if f.dtype != np.uint8 or f.ndim != 3 or f.shape[2] != 3:
            raise BadArgError(OF_E_BADARG, "frame must be an (H, W, 3) uint8 array")
        lp = OfLiveParams()
        lib().of_live_params_default(C.byref(lp))
        for k, val in kw.items():
            if not hasattr(lp, k):
                raise BadArgError(OF_E_BADARG, "unknown live parameter %s" % k)
            setattr(lp, k, val)
        stats = np.zeros(1, STATS_DTYPE)
        u = v = None
        if want_flow:
            u = np.empty((lp.work_height, lp.work_width), np.float64)
            v = np.empty_like(u)
        stage = C.c_int(-1)
        self._check(lib().of_live_step_u8(self._h, C.byref(lp), f.ctypes.data, f.shape[1], f.shape[0], f.strides[0],
                                          stats.ctypes.data, u.ctypes.data if want_flow else None,
                                          v.ctypes.data if want_flow else None, lp.work_width * 8, C.byref(stage)))
        return stage.value, (stats[0] if stage.value == 2 else None), u, v

    def bgr2gray_dev(self, bgr, gray, coeff_bits: int = 14, stream=None):
        """cvtColor(CV_BGR2GRAY) :791 on device tensors: bgr (H, W, 3) uint8 -> gray (H, W) uint8."""
        h, w = gray.shape
        s = _stream_handle(stream, bgr.device)
        self._check(lib().of_bgr2gray_u8_dev(self._h, bgr.data_ptr(), w, h, bgr.stride(0), gray.data_ptr(),
                                             gray.stride(0), coeff_bits, s))

    def live_diff_dev(self, cur_gray, prev_gray, out, learning_rate: float = 0.3, diff_thresh: int = 10,
                      morph_iterations: int = 2, stream=None):
        """:800-813 on device tensors: two (H, W) uint8 gray frames -> out (H, W) float64."""
        h, w = cur_gray.shape
        if prev_gray.stride(0) != cur_gray.stride(0):
            raise BadArgError(OF_E_BADARG, "cur_gray and prev_gray must share one row step")
        s = _stream_handle(stream, cur_gray.device)
        self._check(lib().of_live_diff_f64_dev(self._h, cur_gray.data_ptr(), prev_gray.data_ptr(), w, h,
                                               cur_gray.stride(0), learning_rate, diff_thresh,
                                               morph_iterations, out.data_ptr(), out.stride(0) * 8, s))


class MultiGpu:
    """``of_mgpu``: the frame stream of the live loop sharded across several GPUs from ONE process (one
    context and one host thread per device inside the C library; contiguous blocks of pairs)."""

    def __init__(self, devices=None, win_radius: int = 1, det_eps: float = 0.0, mag_thresh: float = 20.0,
                 count_thresh: int = 500):
        L = lib()
        p = default_params()
        p.win_radius, p.det_eps = int(win_radius), float(det_eps)
        p.mag_thresh, p.count_thresh = float(mag_thresh), int(count_thresh)
        self._h = C.c_void_p()
        if devices is None:
            st = L.of_mgpu_create(None, 0, C.byref(p), C.byref(self._h))
        else:
            arr = (C.c_int * len(devices))(*[int(d) for d in devices])
            st = L.of_mgpu_create(arr, len(devices), C.byref(p), C.byref(self._h))
        if st != OF_OK:
            raise OfError(st, (L.of_mgpu_last_error_string(None) or b"").decode())

    @property
    def device_count(self) -> int:
        return int(lib().of_mgpu_device_count(self._h))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().of_mgpu_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @staticmethod
    def partition(n_pairs: int, n_devices: int, index: int):
        a, b = C.c_int(0), C.c_int(0)
        st = lib().of_mgpu_partition(int(n_pairs), int(n_devices), int(index), C.byref(a), C.byref(b))
        if st != OF_OK:
            raise BadArgError(st, "bad partition arguments")
        return a.value, b.value

    def flow_stream(self, frames, nLevels: int, want_flow: bool = True, want_stats: bool = True, out_u=None,
                    out_v=None, alpha: float = 1.0 / 255, beta: float = 0.0):
        """Same contract as ``Context.flow_stream`` with the pairs split across the devices."""
        f = np.asarray(frames)
        if f.dtype not in (np.float64, np.uint8) or f.ndim != 3 or not f.flags.c_contiguous:
            raise BadArgError(OF_E_BADARG, "frames must be a C-contiguous (n, H, W) float64 or uint8 array")
        n, h, w = f.shape
        u = v = None
        if want_flow:
            u = out_u if out_u is not None else np.empty((n - 1, h, w), np.float64)
            v = out_v if out_v is not None else np.empty((n - 1, h, w), np.float64)
        stats = np.zeros(n - 1, STATS_DTYPE) if want_stats else None
        up = u.ctypes.data if want_flow else None
        vp_ = v.ctypes.data if want_flow else None
        sp = stats.ctypes.data if want_stats else None
        if f.dtype == np.uint8:
            st = lib().of_mgpu_ctf_u8_stream(self._h, f.ctypes.data, n, h * w, w, h, w, float(alpha), float(beta),
                                             int(nLevels), up, vp_, h * w * 8, w * 8, sp)
        else:
            st = lib().of_mgpu_ctf_f64_stream(self._h, f.ctypes.data, n, h * w * 8, w, h, w * 8, int(nLevels), up, vp_,
                                              h * w * 8, w * 8, sp)
        if st != OF_OK:
            msg = (lib().of_mgpu_last_error_string(self._h) or b"").decode()
            raise (NotDivisibleError if st == OF_E_NOT_DIVISIBLE else BadArgError if st == OF_E_BADARG else OfError)(st, msg)
        return u, v, stats
