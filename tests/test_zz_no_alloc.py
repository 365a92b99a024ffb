"""msda_b200_backward_static_scratch (include/msda_b200.h): the 16-bit backward with its scratch in a static device
array of the library instead of a caller-allocated workspace.  ``-m gpu``; collected last on purpose.

Called through the C ABI with ctypes, next to the existing entry point msda_b200_backward (workspace allocated by
the test) on the same inputs.  Both run the same kernels, so they may differ only by the order of the fp32
reductions: grad_sampling_loc / grad_attn_weight to fp32 rounding (2e-5 of max|ref|), grad_value to 2.5 units in the
last place of the 16-bit output -- the tolerances tests/test_dense_gpu.py uses between two runs of the backward.
Two calls in a row with different shapes and inputs: the second finds the first one's contents in the array.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


def _rel(a, ref):
    a = a.double().cpu().numpy()
    ref = ref.double().cpu().numpy()
    return float(np.abs(a - ref).max() / max(np.abs(ref).max(), 1e-30))


def _case(levels, N, M, D, P, dtype, seed, Lq=None):
    """Encoder form (Lq None: the queries are the pyramid's pixels, samples near the query) or a dense decoder form
    (Lq queries anywhere)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    L = len(levels)
    S = sum(h * w for h, w in levels)
    value = torch.randn(N, S, M, D, generator=g).to(dtype)
    if Lq is None:
        pts = []
        for h, w in levels:
            ys = (torch.arange(h, dtype=torch.float32) + 0.5) / h
            xs = (torch.arange(w, dtype=torch.float32) + 0.5) / w
            yy, xx = torch.meshgrid(ys, xs, indexing="ij")
            pts.append(torch.stack([xx.reshape(-1), yy.reshape(-1)], -1))
        ref = torch.cat(pts, 0)
        wh = torch.as_tensor([[w, h] for h, w in levels], dtype=torch.float32)
        off = torch.randn(N, S, M, L, P, 2, generator=g) + torch.arange(1, P + 1).view(1, 1, 1, 1, P, 1) * 0.7
        loc = ref[None, :, None, None, None, :] + off / wh.view(1, 1, 1, L, 1, 2)
        Lq = S
    else:
        loc = torch.rand(N, Lq, M, L, P, 2, generator=g) * 1.1 - 0.05
    attn = torch.softmax(torch.randn(N, Lq, M, L * P, generator=g), -1).view(N, Lq, M, L, P)
    go = torch.randn(N, Lq, M * D, generator=g).to(dtype)
    shapes = torch.as_tensor(levels, dtype=torch.int64)
    return [t.contiguous().to(_dev()) for t in (value, shapes, loc, attn, go)]


def _call(lib, static, value, shapes, loc, attn, go):
    """One backward through the C ABI; static: the new entry point, else msda_b200_backward + a torch workspace."""
    from kinet_b200 import _lib
    N, S, M, D = value.shape
    Lq, L, P = loc.shape[1], loc.shape[3], loc.shape[4]
    dims = _lib.MsdaDims(N, S, M, D, L, Lq, P)
    vcode, acode = _lib.dtype_code(value.dtype), _lib.dtype_code(loc.dtype)
    gv = torch.full_like(value, float("nan"))             # "need not be initialised"
    gl = torch.full_like(loc, float("nan"))
    ga = torch.full_like(attn, float("nan"))
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    common = (value.data_ptr(), vcode, shapes.data_ptr(), None, loc.data_ptr(), attn.data_ptr(), acode, go.data_ptr(),
              gv.data_ptr(), gl.data_ptr(), ga.data_ptr())
    need = int(lib.msda_b200_backward_workspace_bytes(vcode, ctypes.byref(dims)))
    if static:
        rc = lib.msda_b200_backward_static_scratch(*common, ctypes.byref(dims), stream)
        ws = None
    else:
        ws = torch.empty(max(need, 16), dtype=torch.uint8, device=value.device)
        rc = lib.msda_b200_backward(*common, ws.data_ptr(), need, ctypes.byref(dims), stream)
    assert rc == 0, lib.msda_b200_error_string(rc)
    launches = lib.msda_b200_last_launch_count()
    torch.cuda.synchronize()
    return (gv, gl, ga), launches, need


def _compare(a, b, dtype):
    ulp = 2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -11
    for t in a + b:
        assert torch.isfinite(t.float()).all()
    assert _rel(a[0], b[0]) <= 2.5 * ulp
    assert _rel(a[1], b[1]) <= 2e-5
    assert _rel(a[2], b[2]) <= 2e-5


def test_static_scratch_backward_matches_workspace_backward_two_calls_in_a_row():
    from kinet_b200 import _lib
    lib = _lib.load()
    # first call: encoder form, bf16, D = 32 -- the dense-window kernel + the per-row kernel + the rounding pass
    # (accumulator, work counter and flags all live in the array); the larger of the two shapes
    first = _case([(40, 62), (20, 31), (10, 16), (5, 8)], 2, 4, 32, 4, torch.bfloat16, seed=1)
    # second call right after it: another dtype, a smaller pyramid, a dense decoder form (per-row kernel with the fp32
    # accumulator in the array, which now holds the first call's sums, counter and flags)
    second = _case([(24, 36), (12, 18), (6, 9)], 1, 8, 32, 4, torch.float16, seed=2, Lq=700)
    third = _case([(40, 62), (20, 31), (10, 16), (5, 8)], 2, 4, 32, 4, torch.bfloat16, seed=3)

    got1, launches1, need1 = _call(lib, True, *first)
    got2, launches2, need2 = _call(lib, True, *second)
    got3, _, _ = _call(lib, True, *third)
    assert 0 < need2 < need1 <= _lib.STATIC_SCRATCH_BYTES
    assert launches1 == 3 and launches2 == 2, (launches1, launches2)

    want1, base1, _ = _call(lib, False, *first)
    want2, base2, _ = _call(lib, False, *second)
    want3, _, _ = _call(lib, False, *third)
    assert (base1, base2) == (launches1, launches2)
    _compare(got1, want1, torch.bfloat16)
    _compare(got2, want2, torch.float16)
    _compare(got3, want3, torch.bfloat16)
    # and once more after the workspace calls: the array is still usable, the result still the same
    again, _, _ = _call(lib, True, *first)
    _compare(again, want1, torch.bfloat16)


def test_static_scratch_rejects_sizes_beyond_the_array():
    from kinet_b200 import _lib
    lib = _lib.load()
    dev = _dev()
    # BASELINE.json configs[1] (N=2, S=22223, M=8, D=32, bf16, encoder form) needs 91 MB of scratch
    dims = _lib.MsdaDims(2, 22223, 8, 32, 4, 22223, 4)
    assert lib.msda_b200_backward_workspace_bytes(_lib.MSDA_BF16, ctypes.byref(dims)) > _lib.STATIC_SCRATCH_BYTES
    one = torch.zeros(64, device=dev)
    p = one.data_ptr()
    rc = lib.msda_b200_backward_static_scratch(p, _lib.MSDA_BF16, p, None, p, p, _lib.MSDA_F32, p, p, p, p,
                                               ctypes.byref(dims), None)
    assert rc == _lib.MSDA_E_SCRATCH_LIMIT
    assert lib.msda_b200_last_launch_count() == 0
    msg = lib.msda_b200_error_string(rc).decode()
    assert "MSDA_B200_STATIC_SCRATCH_BYTES" in msg and "msda_b200_backward" in msg
    torch.cuda.synchronize()


def test_static_scratch_backward_allocates_nothing():
    """Stream capture forbids cudaMalloc / cudaFree (they fail and invalidate the capture): the call records into a
    CUDA graph, and the replayed graph gives the same result."""
    from kinet_b200 import _lib
    lib = _lib.load()
    case = _case([(20, 31), (10, 16), (5, 8), (3, 4)], 2, 2, 32, 4, torch.bfloat16, seed=4)
    want, _, _ = _call(lib, False, *case)
    _call(lib, True, *case)                                # the array's address is resolved outside the capture
    value, shapes, loc, attn, go = case
    N, S, M, D = value.shape
    dims = _lib.MsdaDims(N, S, M, D, loc.shape[3], loc.shape[1], loc.shape[4])
    gv, gl, ga = torch.empty_like(value), torch.empty_like(loc), torch.empty_like(attn)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        rc = lib.msda_b200_backward_static_scratch(
            value.data_ptr(), _lib.MSDA_BF16, shapes.data_ptr(), None, loc.data_ptr(), attn.data_ptr(), _lib.MSDA_F32,
            go.data_ptr(), gv.data_ptr(), gl.data_ptr(), ga.data_ptr(), ctypes.byref(dims),
            ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.msda_b200_error_string(rc)
    for _ in range(2):
        gv.fill_(float("nan"))
        graph.replay()
        torch.cuda.synchronize()
        _compare((gv, gl, ga), want, torch.bfloat16)
