"""GPU: ctrhot_mlp_f32_static -- the fp32 tower with its scratch in a file-scope device array (no allocation anywhere) --
against ctrhot_mlp_f32 (caller workspace) through the C-ABI, on two calls in a row, and against the float64 oracle.

Collected LAST on purpose.  The entry point has no counterpart in the reference
(include/ctrhot.h).  Passes alone in a fresh process and after the other tests:
it depends on no state but the library's own static array, which it assumes nothing about.
"""
import ctypes

import numpy as np
import pytest
import torch

from oracle import ctr_oracle as O
from recommender_system_b200 import _lib

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 16 << 20      # CTRHOT_MLP_STATIC_SCRATCH_BYTES in include/ctrhot.h
FP32_TOL = 2e-5               # the fp32 tower's tolerance in tests/test_gpu_parity.py::test_dnn_fp32_and_tf32


def _case(seed, M, K, hidden, act):
    rng = np.random.default_rng(seed)
    dims = [K] + list(hidden) + [1]
    x = rng.uniform(-0.5, 1.0, (M, K)).astype(np.float32)
    ws = []
    for i in range(len(dims) - 1):
        lim = np.sqrt(6.0 / (dims[i] + dims[i + 1]))
        ws.append((rng.uniform(-lim, lim, (dims[i], dims[i + 1])).astype(np.float32),
                   rng.normal(0, 0.1, (dims[i + 1],)).astype(np.float32)))
    return x, ws, act


def _descs(ws, act, dev):
    """ctrhot_dense_desc array for hidden layers with `act` and a linear output layer; returns (array, keep-alive list)."""
    keep = []
    arr = (_lib.DenseDesc * len(ws))()
    for i, (w, b) in enumerate(ws):
        wt, bt = torch.from_numpy(w).to(dev).contiguous(), torch.from_numpy(b).to(dev).contiguous()
        keep += [wt, bt]
        arr[i].weight, arr[i].bias, arr[i].act_param, arr[i].packed = wt.data_ptr(), bt.data_ptr(), None, None
        arr[i].in_dim, arr[i].out_dim = w.shape
        arr[i].act = _lib.ACT_BY_NAME[act] if i + 1 < len(ws) else _lib.ACT_NONE
        arr[i].alpha_rows, arr[i].dice_eps, arr[i].reserved = 1, 0.0, 0
    return arr, keep


def _run_both(lib, x, ws, act, dev):
    M = x.shape[0]
    xd = torch.from_numpy(x).to(dev)
    arr, keep = _descs(ws, act, dev)
    stream = torch.cuda.current_stream().cuda_stream
    # existing entry point: the caller (here: the test, through torch) owns the workspace
    nbytes = lib.ctrhot_mlp_workspace_bytes(M, arr, len(ws))
    assert nbytes > 0
    work = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    want = torch.full((M, 1), float("nan"), dtype=torch.float32, device=dev)
    _lib.check(lib.ctrhot_mlp_f32(xd.data_ptr(), xd.stride(0), M, arr, len(ws), want.data_ptr(), 1, work.data_ptr(), nbytes,
                                  stream), "ctrhot_mlp_f32")
    # new entry point: inputs and outputs only
    got = torch.full((M, 1), float("nan"), dtype=torch.float32, device=dev)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    _lib.check(lib.ctrhot_mlp_f32_static(xd.data_ptr(), xd.stride(0), M, arr, len(ws), got.data_ptr(), 1, stream),
               "ctrhot_mlp_f32_static")
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    del keep
    return got.cpu().numpy(), want.cpu().numpy(), free0 - free1


def test_static_scratch_tower_matches_workspace_tower_twice_in_a_row(lib):
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    # two calls in a row that leave different leftovers in the shared array: a large tower first, then a smaller one with
    # other data, widths and row count (its buffers overlap the first call's in a different layout), then the first again
    cases = [_case(11, 4099, 221, [256, 128, 64], "relu"), _case(12, 300, 77, [40, 24], "tanh"),
             _case(11, 4099, 221, [256, 128, 64], "relu")]
    leaked = []
    for i, (x, ws, act) in enumerate(cases):
        got, want, d_free = _run_both(lib, x, ws, act, dev)
        assert np.isfinite(got).all(), "call %d: output not fully written" % i
        # same kernels on the same inputs: identical bits, not merely close
        assert np.array_equal(got, want), "call %d: max |static - workspace| = %.3g" % (i, np.max(np.abs(got - want)))
        ref = O.dnn_layer(x.astype(np.float64), [(w.astype(np.float64), b.astype(np.float64)) for w, b in ws[:-1]],
                          tuple(a.astype(np.float64) for a in ws[-1]), act)
        err = float(np.max(np.abs(got.astype(np.float64) - ref)))
        assert err <= FP32_TOL, "call %d: max |static - oracle| = %.3g > %.3g" % (i, err, FP32_TOL)
        leaked.append(d_free)
    # no allocation on any call but possibly the first of the process (the symbol's address is resolved once per device;
    # lazy module loading may take device memory then)
    assert leaked[1] == 0 and leaked[2] == 0, "free device memory moved across calls: %r" % (leaked,)


def test_static_scratch_single_layer_and_empty_batch(lib):
    """One layer needs no scratch at all; M = 0 is a no-op."""
    dev = torch.device("cuda", 0)
    x, ws, act = _case(13, 65, 19, [], "relu")
    got, want, _ = _run_both(lib, x, ws, act, dev)
    assert np.array_equal(got, want)
    arr, keep = _descs(ws, act, dev)
    xd = torch.from_numpy(x).to(dev)
    out = torch.zeros((1, 1), dtype=torch.float32, device=dev)
    assert lib.ctrhot_mlp_f32_static(xd.data_ptr(), xd.stride(0), 0, arr, len(ws), out.data_ptr(), 1, None) == _lib.OK


def test_static_scratch_rejects_sizes_beyond_its_maximum(lib):
    dev = torch.device("cuda", 0)
    x, ws, act = _case(14, 8, 32, [256], "relu")
    arr, keep = _descs(ws, act, dev)
    xd = torch.from_numpy(x).to(dev)
    out = torch.zeros((8, 1), dtype=torch.float32, device=dev)
    fits = SCRATCH_BYTES // (2 * 256 * 4)                      # 8 192 rows at 256 hidden units fill the array exactly
    # the size check comes before any launch: the refusal must not touch x / out (which hold 8 rows only)
    rc = lib.ctrhot_mlp_f32_static(xd.data_ptr(), xd.stride(0), fits + 1, arr, len(ws), out.data_ptr(), 1, None)
    assert rc == _lib.ESHAPE
    msg = lib.ctrhot_last_cuda_error().decode()
    assert "static scratch" in msg and str(SCRATCH_BYTES) in msg, msg
    rc = lib.ctrhot_mlp_f32_static(xd.data_ptr(), xd.stride(0), 1 << 40, arr, len(ws), out.data_ptr(), 1, None)
    assert rc == _lib.ESHAPE                                   # no overflow in the size arithmetic either
    torch.cuda.synchronize()
    # exactly at the limit it runs
    xl = torch.rand((fits, 32), dtype=torch.float32, device=dev)
    ol = torch.full((fits, 1), float("nan"), dtype=torch.float32, device=dev)
    _lib.check(lib.ctrhot_mlp_f32_static(xl.data_ptr(), xl.stride(0), fits, arr, len(ws), ol.data_ptr(), 1,
                                         torch.cuda.current_stream().cuda_stream), "ctrhot_mlp_f32_static at the limit")
    torch.cuda.synchronize()
    assert torch.isfinite(ol).all()
