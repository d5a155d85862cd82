"""GPU suite (-m gpu), round 2 additions: the parity holes the round-1 review listed.

  * magnitudes that leave the fast path of the shared-reciprocal division (lk_common.cuh: div2_fast /
    ddiv_slow) and denormal determinants, bit-exact at r = 1 (ParallelVideoPyr.cpp:514,518: divide);
  * BASELINE's full-size configurations with several pairs, including the gesture-like stream whose
    large exactly-zero regions give interior det == 0 (the data regime of the live loop, :800-813);
  * multi-GPU: a stream sharded over devices reproduces the single-GPU rows (statistics :845-888),
    both in one process (of_mgpu_*, a context per device from a thread per device) and with one
    process per GPU (torchrun + NCCL gather).  Skipped on a one-GPU box.
"""
import os
import subprocess
import sys
import threading

import numpy as np
import pytest

from test_gpu_parity import FLOW_TOL, check, ofb_synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def n_gpus():
    import torch
    return torch.cuda.device_count()


# ------------------------------------------------------------------ division paths
@pytest.mark.parametrize("scale", [1e70, 1e76, 1e-30, 1e-70, 1e-77, 1e-79, 1e-80, 1e-81, 1e-100])
def test_extreme_magnitudes_bit_exact(ofb, c_oracle, table_pair, scale):
    """u = nu / det is scale-free, but nu and det are not: at 1e-70 the numerators drop below the
    range the fast division accepts (library path), from 1e-77 on the determinants are denormal
    and at 1e-100 they underflow to 0 (mask 0); 1e70 squares to 1e280.  All bit-exact vs the oracle."""
    a, b = table_pair
    a, b = a[:192, :256] * scale, b[:192, :256] * scale
    with ofb.Context(0) as ctx:
        got = ctx.coarseToFineEstimation(a, b, 3, want_mask=True)
        got1 = ctx.getLucasKanadeOpticalFlow(a, b, want_mask=True)
    want = c_oracle.coarse_to_fine(a, b, 3, want_mask=True)
    if scale == 1e-100:
        assert not want[2].any()
    elif 1e-77 <= scale <= 1e70:
        assert want[2].sum() > 40000
    assert np.isfinite(want[0]).all()
    check(got, want, 1, "scale %g" % scale)
    check(got1, c_oracle.lk_single(a, b, want_mask=True), 1, "single level, scale %g" % scale)


def test_mixed_magnitudes_r7(ofb, c_oracle, table_pair):
    """r = 7: a frame whose left half is scaled by 1e-70 (library division) and whose right half is
    not (fast division) in the same warps."""
    a, b = table_pair
    a, b = a[:128, :256].copy(), b[:128, :256].copy()
    a[:, :128] *= 1e-70
    b[:, :128] *= 1e-70
    with ofb.Context(0, win_radius=7) as ctx:
        got = ctx.getLucasKanadeOpticalFlow(a, b, want_mask=True)
    check(got, c_oracle.lk_single(a, b, 7, want_mask=True), 7, "mixed magnitudes")


# ------------------------------------------------------------------ full-size, several pairs
@pytest.mark.parametrize("cfg", [(1080, 1920, 3, 7), (1080, 1920, 3, 1), (2160, 3840, 4, 7), (2160, 3840, 4, 1)])
@pytest.mark.parametrize("gesture", [False, True])
def test_full_size_streams_vs_oracle(ofb, c_oracle, cfg, gesture):
    """BASELINE configs 3/4 (synthetic 1080p L=3 and 4K L=4): four consecutive pairs of the plain and
    of the gesture-like stream (moving disc, exactly 0 elsewhere) against the C oracle, with the mask."""
    import torch
    h, w, L, r = cfg
    n = 5
    fr = ofb_synth().synth_frames(n, h, w, seed=1234, gesture=gesture)
    d = torch.from_numpy(fr).cuda()
    U = torch.empty((n - 1, h, w), dtype=torch.float64, device="cuda")
    V = torch.empty_like(U)
    M = torch.empty((n - 1, h, w), dtype=torch.uint8, device="cuda")
    S = torch.empty((n - 1, 4), dtype=torch.float64, device="cuda")
    with ofb.Context(0, win_radius=r, mag_thresh=1.0) as ctx:
        ctx.flow_batch_dev(d, L, U, V, S, M)
        torch.cuda.synchronize()
    for k in range(n - 1):
        want = c_oracle.coarse_to_fine(fr[k], fr[k + 1], L, r, want_mask=True)
        if gesture:
            assert (~want[2]).sum() > 0.5 * h * w           # mostly det == 0
        got = (U[k].cpu().numpy(), V[k].cpu().numpy(), M[k].cpu().numpy().astype(bool))
        check(got, want, r, "%s gesture=%s pair %d" % (cfg, gesture, k))
        s = c_oracle.gesture_stats(got[0], got[1], 1.0, 500)
        row = S[k].cpu().numpy()
        assert (row[0], row[1], row[2]) == (s["sum_x"], s["sum_y"], s["count"])


# ------------------------------------------------------------------ streams and contexts
def test_calls_on_different_streams_are_ordered(ofb, c_oracle):
    """All calls on a context share its scratch: a call on another stream waits for the previous one."""
    import torch
    fr = ofb_synth().synth_frames(9, 272, 480, seed=7)
    d = torch.from_numpy(fr).cuda()
    U = torch.empty((8, 272, 480), dtype=torch.float64, device="cuda")
    V = torch.empty_like(U)
    u1 = torch.empty((272, 480), dtype=torch.float64, device="cuda")
    v1 = torch.empty_like(u1)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    torch.cuda.synchronize()
    with ofb.Context(0, win_radius=7) as ctx:
        for _ in range(5):
            ctx.flow_batch_dev(d, 3, U, V, stream=s1)
            ctx.coarseToFineEstimation_dev(d[0], d[1], 3, u1, v1, stream=s2)
        torch.cuda.synchronize()
    want = c_oracle.coarse_to_fine(fr[0], fr[1], 3, 7)
    assert torch.equal(U[0], u1) and torch.equal(V[0], v1)
    assert max(np.abs(u1.cpu().numpy() - want[0]).max(), np.abs(v1.cpu().numpy() - want[1]).max()) <= FLOW_TOL
    want7 = c_oracle.coarse_to_fine(fr[7], fr[8], 3, 7)
    assert np.abs(U[7].cpu().numpy() - want7[0]).max() <= FLOW_TOL


def test_repeated_small_calls_replay_a_cuda_graph(ofb, c_oracle, table_pair, monkeypatch):
    """The reference's call pattern (one pair per call, same buffers): from the third identical call on the
    launch sequence is a graph replay; results stay those of the plain launches; new pointers, new
    parameters or a larger scratch arena never replay a stale graph."""
    import torch
    a, b = table_pair
    d = torch.from_numpy(np.stack([a, b])).cuda()
    u = torch.empty((448, 640), dtype=torch.float64, device="cuda")
    v = torch.empty_like(u)
    want = c_oracle.coarse_to_fine(a, b, 7)
    with ofb.Context(0) as ctx:
        n0 = ctx.launch_count
        for k in range(5):
            u.zero_()
            ctx.coarseToFineEstimation_dev(d[0], d[1], 7, u, v)
            torch.cuda.synchronize()
            assert np.array_equal(u.cpu().numpy(), want[0]) and np.array_equal(v.cpu().numpy(), want[1]), k
        assert ctx.graph_replays == 4                       # calls 2..5 (the second one captures, then replays)
        per_call = (ctx.launch_count - n0) // 5
        assert per_call >= 13 and (ctx.launch_count - n0) == 5 * per_call
        u2 = torch.empty_like(u)                            # other pointers: plain launches again
        ctx.coarseToFineEstimation_dev(d[0], d[1], 7, u2, v)
        assert ctx.graph_replays == 4
        ctx.set_params(win_radius=2)                        # other parameters
        ctx.coarseToFineEstimation_dev(d[0], d[1], 7, u, v)
        assert ctx.graph_replays == 4
        ctx.set_params(win_radius=1)
        big = torch.zeros((2, 1024, 2048), dtype=torch.float64, device="cuda")      # grows the arena: graphs dropped
        ub = torch.empty((1024, 2048), dtype=torch.float64, device="cuda")
        ctx.coarseToFineEstimation_dev(big[0], big[1], 3, ub, ub.clone())
        for k in range(3):
            ctx.coarseToFineEstimation_dev(d[0], d[1], 7, u, v)
        torch.cuda.synchronize()
        assert ctx.graph_replays == 6
        assert np.array_equal(u.cpu().numpy(), want[0])
    monkeypatch.setenv("OF_B200_GRAPH", "0")
    with ofb.Context(0) as ctx:
        for k in range(3):
            ctx.coarseToFineEstimation_dev(d[0], d[1], 7, u, v)
        assert ctx.graph_replays == 0


def test_write_back_rejected_where_inputs_are_const(ofb):
    import torch
    z = torch.zeros((2, 64, 64), dtype=torch.float64, device="cuda")
    u = torch.empty((64, 64), dtype=torch.float64, device="cuda")
    with ofb.Context(0, write_back_warped=True) as ctx:
        with pytest.raises(ofb.BadArgError):
            ctx.coarseToFineEstimation_dev(z[0], z[1], 2, u, u.clone())
        with pytest.raises(ofb.BadArgError):
            ctx.flow_stream(np.zeros((3, 64, 64)), 2)
        ctx.coarseToFineEstimation_dev(z[0], z[1], 1, u, u.clone())      # one level: nothing is warped


def test_bad_parameters_rejected_at_create(ofb):
    for kw in ({"win_radius": 0}, {"win_radius": 16}, {"det_eps": float("nan")}, {"det_eps": -1.0}):
        with pytest.raises(ofb.OfError):
            ofb.Context(0, **kw)


def test_stream_last_frame_not_overread(ofb, c_oracle):
    """W * H * 8 not a multiple of 256: the caller's buffers end right after the last frame / plane."""
    h, w, n = 27, 48, 4                                       # 27 * 48 * 8 = 10368 = 40.5 * 256
    fr = ofb_synth().synth_frames(n, h, w, seed=3)
    guard = 1024
    fin = np.empty(n * h * w + guard)
    fin[:n * h * w] = fr.ravel()
    fin[n * h * w:] = np.nan
    out = np.full(2 * ((n - 1) * h * w + guard), 7.0)
    u = out[:(n - 1) * h * w].reshape(n - 1, h, w)
    v = out[(n - 1) * h * w + guard:2 * (n - 1) * h * w + guard].reshape(n - 1, h, w)
    with ofb.Context(0) as ctx:
        ctx.flow_stream(fin[:n * h * w].reshape(n, h, w), 1, out_u=u, out_v=v, want_stats=False)
    assert (out[(n - 1) * h * w:(n - 1) * h * w + guard] == 7.0).all() and (out[2 * (n - 1) * h * w + guard:] == 7.0).all()
    for k in range(n - 1):
        uo, vo = c_oracle.lk_single(fr[k], fr[k + 1])
        assert np.array_equal(u[k], uo) and np.array_equal(v[k], vo)


# ------------------------------------------------------------------ multi-GPU
def test_mgpu_entry_point_matches_single_gpu(ofb, c_oracle):
    """of_mgpu_ctf_f64_stream / of_mgpu_ctf_u8_stream over every visible GPU == one GPU, row for row."""
    fr = ofb_synth().synth_frames(12, 96, 160, seed=3, gesture=True)
    raw = np.clip(np.rint(fr * 255), 0, 255).astype(np.uint8)
    for r in (1, 7):
        with ofb.Context(0, win_radius=r, mag_thresh=0.5, count_thresh=100) as ctx:
            U, V, S = ctx.flow_stream(fr, 3)
            U8, V8, S8 = ctx.flow_stream(raw, 3)
        with ofb.MultiGpu(None, win_radius=r, mag_thresh=0.5, count_thresh=100) as m:
            assert m.device_count == n_gpus()
            Um, Vm, Sm = m.flow_stream(fr, 3)
            Um8, Vm8, Sm8 = m.flow_stream(raw, 3)
            _, _, So = m.flow_stream(fr, 3, want_flow=False)
        assert np.array_equal(U, Um) and np.array_equal(V, Vm) and np.array_equal(S, Sm) and np.array_equal(S, So)
        assert np.array_equal(U8, Um8) and np.array_equal(V8, Vm8) and np.array_equal(S8, Sm8)
        want = c_oracle.coarse_to_fine(fr[5], fr[6], 3, r)
        assert np.abs(Um[5] - want[0]).max() <= (0.0 if r == 1 else FLOW_TOL)
    with ofb.MultiGpu([0]) as m:                              # more devices than pairs / explicit list
        u, v, s = m.flow_stream(fr[:2], 1)
        assert u.shape == (1, 96, 160)
    with pytest.raises(ofb.OfError):
        ofb.MultiGpu([0, 0])


def test_mgpu_partition(ofb):
    for n in (0, 1, 7, 194):
        for nd in (1, 2, 3, 8):
            got = []
            for i in range(nd):
                a, b = ofb.MultiGpu.partition(n, nd, i)
                got += list(range(a, b))
            assert got == list(range(n))


def test_two_devices_in_one_process_from_two_threads(ofb, c_oracle):
    """Distinct contexts on distinct devices from distinct threads of one process (of_b200.h's
    threading model; the per-device shared-memory opt-in of the level kernels)."""
    if n_gpus() < 2:
        pytest.skip("needs two GPUs")
    fr = ofb_synth().synth_frames(3, 272, 480, seed=21)
    res, errs = {}, []

    def work(dev, r):
        try:
            with ofb.Context(dev, win_radius=r) as ctx:
                for _ in range(3):
                    res[dev] = ctx.coarseToFineEstimation(fr[dev], fr[dev + 1], 3, want_mask=True)
        except Exception as e:                                 # pragma: no cover
            errs.append(e)

    th = [threading.Thread(target=work, args=(1, 7)), threading.Thread(target=work, args=(0, 1))]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs, errs
    check(res[0], c_oracle.coarse_to_fine(fr[0], fr[1], 3, 1, want_mask=True), 1, "device 0")
    check(res[1], c_oracle.coarse_to_fine(fr[1], fr[2], 3, 7, want_mask=True), 7, "device 1")


def test_cpp_caller_drives_every_gpu(ofb, tmp_path):
    """tests/shim_check.cpp --mgpu: a C++ caller of the drop-in (what the reference is) runs the stream on
    every GPU of the box through of_mgpu_ctf_f64_stream and compares with one GPU."""
    from test_shim import build_shim_check
    exe = build_shim_check(ofb, tmp_path)
    fr = ofb_synth().synth_frames(9, 96, 160, seed=5, gesture=True)
    p = str(tmp_path / "frames.f64")
    fr.tofile(p)
    for r in ("1", "7"):
        out = subprocess.run([exe, "--mgpu", "160", "96", "9", p, "3", r], capture_output=True, text=True)
        assert out.returncode == 0, out.stderr
        assert out.stdout.strip() == "mgpu ok %d" % n_gpus()


def test_torchrun_two_ranks_reproduce_single_gpu_rows(ofb, tmp_path):
    """One process per GPU (the bench's launch): a 33-frame stream split with partition_pairs, per-pair
    statistics gathered to rank 0 over NCCL == the single-GPU flow_stream rows."""
    if n_gpus() < 2:
        pytest.skip("needs two GPUs")
    out = str(tmp_path / "rows.npy")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29611", os.path.join(ROOT, "tests", "shard_worker.py"), out]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    rows = np.load(out)
    fr = ofb_synth().synth_frames(33, 272, 480, seed=1234, gesture=True)
    with ofb.Context(0, win_radius=7, mag_thresh=0.5, count_thresh=100) as ctx:
        _, _, S = ctx.flow_stream(fr, 3, want_flow=False)
    ref = np.stack([S["sum_x"], S["sum_y"], S["count"], S["detected"]], axis=1)
    assert rows.shape == (32, 4) and np.array_equal(rows, ref)
