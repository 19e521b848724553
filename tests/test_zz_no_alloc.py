"""fvp_nms_topk_static: the allocation-free variant of fvp_nms_topk (workspace in a static device array of
FVP_NMS_STATIC_MAX_CELLS floats inside the library), called through the C ABI and compared with fvp_nms_topk.

The entry point has no counterpart in the reference, so the existing
entry point -- itself pinned bit for bit to the oracle and the golden vectors by tests/test_gpu_parity.py -- is the
checker.  Tolerance: none; top-k indices and values are bit-exact in this suite.

Named test_zz_* so that pytest collects it last: a failure here cannot hide another test under `pytest -x`.
Passes alone in a fresh process (it performs the per-device set-up itself) and after the rest of the suite.
"""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from faster_voxelpose_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEV = "cuda:0"


def _header_limit() -> int:
    header = open(os.path.join(ROOT, "include", "fvp_b200.h")).read()
    m = re.search(r"#define\s+FVP_NMS_STATIC_MAX_CELLS\s+\(1\s*<<\s*(\d+)\)", header)
    assert m, "FVP_NMS_STATIC_MAX_CELLS not found in include/fvp_b200.h"
    return 1 << int(m.group(1))


def _maps(seed, B, X, Y):
    rng = np.random.default_rng(seed)
    pm = rng.standard_normal((B, X, Y)).astype(np.float32)
    pm[:, ::3, ::2] = np.round(pm[:, ::3, ::2], 1)      # plateaus and exact ties: the tie rule is exercised
    if X > 16:
        pm[0, 10:14, 10:14] = 2.5
    return torch.from_numpy(pm).to(DEV)


def _outputs(B, P):
    return (torch.full((B, P), -7.0, dtype=torch.float32, device=DEV),
            torch.full((B, P, 2), -7, dtype=torch.int64, device=DEV),
            torch.full((B, P), -7, dtype=torch.int64, device=DEV))


def _existing(lib, hm, P):
    B, X, Y = hm.shape
    scratch = torch.empty((B, X * Y), dtype=torch.float32, device=DEV)
    vals, idx2, flat = _outputs(B, P)
    L.check(lib.fvp_nms_topk(L.ptr(hm), B, X, Y, P, L.ptr(scratch), L.ptr(vals), L.ptr(idx2), L.ptr(flat),
                             L.stream_handle(hm.device)), "fvp_nms_topk")
    return vals, idx2, flat


def test_limit_is_declared_and_oversize_is_rejected_clearly():
    """No GPU needed: the size check precedes every CUDA call (the pointers are never dereferenced on the host)."""
    lib = L.load()
    limit = _header_limit()
    assert limit == L.NMS_STATIC_MAX_CELLS == 1 << 20
    dummy = (ctypes.c_float * 4)()
    a = ctypes.addressof(dummy)
    B, X, Y = 17, 256, 256                              # 17 * 65536 > 2^20
    assert B * X * Y > limit
    rc = lib.fvp_nms_topk_static(a, B, X, Y, 10, a, a, a, None)
    msg = lib.fvp_last_error().decode()
    assert rc == -1, rc
    assert "exceeds" in msg and str(limit) in msg and "fvp_nms_topk" in msg, msg
    rc = lib.fvp_nms_topk_static(a, 40000, 40000, 40000, 10, a, a, a, None)      # product overflows 32 bits
    assert rc == -1 and "exceeds" in lib.fvp_last_error().decode()
    rc = lib.fvp_nms_topk_static(None, 1, 8, 8, 4, a, a, a, None)
    assert rc == -1 and "null" in lib.fvp_last_error().decode()
    rc = lib.fvp_nms_topk_static(a, 1, 3, 3, 10, a, a, a, None)
    assert rc == -1 and "P" in lib.fvp_last_error().decode()


@pytest.mark.gpu
def test_static_workspace_matches_fvp_nms_topk_on_two_calls_in_a_row():
    lib = L.load()
    limit = _header_limit()
    # first call fills the whole workspace (16 x 256 x 256 = the limit); the second is smaller and of another shape, so
    # it runs over what the first left behind; the third returns to the first input (same answer expected again)
    cases = [(11, 16, 256, 256, 32), (12, 4, 80, 80, 10), (11, 16, 256, 256, 32), (13, 3, 7, 7, 5), (14, 1, 3, 3, 9)]
    assert cases[0][1] * cases[0][2] * cases[0][3] == limit
    inputs = [(_maps(seed, B, X, Y), P) for seed, B, X, Y, P in cases]
    expect = [_existing(lib, hm, P) for hm, P in inputs]
    torch.cuda.synchronize()

    # set-up call (looks the symbol address up on this device), then measure that later calls allocate nothing
    hm0, P0 = inputs[1]
    warm = _outputs(hm0.shape[0], P0)
    L.check(lib.fvp_nms_topk_static(L.ptr(hm0), *hm0.shape, P0, L.ptr(warm[0]), L.ptr(warm[1]), L.ptr(warm[2]),
                                    L.stream_handle(hm0.device)), "fvp_nms_topk_static")
    torch.cuda.synchronize()
    outs = [_outputs(hm.shape[0], P) for hm, P in inputs]
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info(0)
    torch_before = torch.cuda.memory_reserved(0)
    for (hm, P), (vals, idx2, flat) in zip(inputs, outs):       # back to back on one stream, no sync in between
        B, X, Y = hm.shape
        L.check(lib.fvp_nms_topk_static(L.ptr(hm), B, X, Y, P, L.ptr(vals), L.ptr(idx2), L.ptr(flat),
                                        L.stream_handle(hm.device)), "fvp_nms_topk_static")
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info(0)
    assert torch.cuda.memory_reserved(0) == torch_before
    assert free_after >= free_before, (free_before, free_after)  # no device memory taken by the calls

    for i, ((vals, idx2, flat), (ev, ei, ef)) in enumerate(zip(outs, expect)):
        assert torch.equal(flat, ef), f"call {i}: flat indices differ"
        assert torch.equal(idx2, ei), f"call {i}: 2-D indices differ"
        assert torch.equal(vals.view(torch.int32), ev.view(torch.int32)), f"call {i}: values differ"
    assert torch.equal(outs[0][2], outs[2][2])          # same input, same answer, whatever ran in between


@pytest.mark.gpu
def test_static_workspace_equals_the_oracle(oracle_lib):
    """Belt and braces: the new entry point against the CPU oracle directly (ties, negatives, NaN-free maps)."""
    lib = L.load()
    for seed, B, X, Y, P in [(21, 2, 80, 80, 10), (22, 5, 24, 24, 10)]:
        hm = _maps(seed, B, X, Y)
        vals, idx2, flat = _outputs(B, P)
        L.check(lib.fvp_nms_topk_static(L.ptr(hm), B, X, Y, P, L.ptr(vals), L.ptr(idx2), L.ptr(flat),
                                        L.stream_handle(hm.device)), "fvp_nms_topk_static")
        rv, ri, rf = oracle_lib.nms_topk(hm.cpu().numpy(), P)
        assert np.array_equal(flat.cpu().numpy(), rf)
        assert np.array_equal(idx2.cpu().numpy(), ri)
        assert np.array_equal(vals.cpu().numpy(), rv)
