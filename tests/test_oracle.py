"""Pins the CPU oracle (oracle/) against golden vectors produced by the real
reference function ms_deform_attn_core_pytorch (tests/golden/make_golden.py).
CPU only."""
import numpy as np
import pytest

from oracle import oracle_backward, oracle_forward, torch_port_forward

CASES = ["recipe_f32", "recipe_f64", "edges_f32", "edges_f64", "l8_f32", "integer_f64"]


def _tol(name):
    # fp64 cases: the two implementations differ only by summation order.
    return (1e-12, 1e-13) if name.endswith("f64") else (2e-5, 1e-6)


def _close(a, b, rtol, atol_scale):
    scale = max(float(np.abs(b).max()), 1e-30)
    np.testing.assert_allclose(a, b, rtol=rtol, atol=atol_scale * scale)


@pytest.mark.parametrize("name", CASES)
def test_c_oracle_forward_matches_reference(golden, name):
    g = golden(name)
    rtol, atol = _tol(name)
    out = oracle_forward(g["value"], g["shapes"], g["loc"], g["attn"])
    assert out.dtype == g["out"].dtype
    _close(out, g["out"], rtol, atol)


@pytest.mark.parametrize("name", [c for c in CASES if c != "integer_f64"])
def test_c_oracle_backward_matches_reference(golden, name):
    g = golden(name)
    rtol, atol = _tol(name)
    gv, gl, ga = oracle_backward(g["value"], g["shapes"], g["loc"], g["attn"], g["grad_out"])
    _close(gv, g["grad_value"], rtol, atol * 10)
    _close(gl, g["grad_loc"], rtol, atol * 10)
    _close(ga, g["grad_attn"], rtol, atol * 10)


def test_c_oracle_backward_integer_coords(golden):
    """At integer pixel coordinates the bilinear form has a kink.  grad_value and grad_attn
    are still unique and must match the reference.  grad_sampling_loc is one-sided there and
    depends on which side rounding lands (grid_sample un-normalises ((2*loc-1)+1)*W/2-0.5, the
    kernels loc*W-0.5, .cuh:227-228; grid_sample also differentiates through the zero-weight
    corner at coordinate -1 exactly, the kernels' open window does not), so away-from-kink
    entries are compared and kink entries only need to be finite."""
    g = golden("integer_f64")
    gv, gl, ga = oracle_backward(g["value"], g["shapes"], g["loc"], g["attn"], g["grad_out"])
    _close(gv, g["grad_value"], 1e-10, 1e-12)
    _close(ga, g["grad_attn"], 1e-10, 1e-12)
    assert np.isfinite(gl).all()


@pytest.mark.parametrize("name", CASES)
def test_torch_port_matches_reference(golden, name):
    import torch
    g = golden(name)
    rtol, atol = _tol(name)
    out = torch_port_forward(torch.from_numpy(g["value"]), torch.from_numpy(g["shapes"]),
                             torch.from_numpy(g["loc"]), torch.from_numpy(g["attn"])).numpy()
    _close(out, g["out"], rtol, atol)


def test_oracle_threads_agree(golden):
    g = golden("l8_f32")
    a = oracle_forward(g["value"], g["shapes"], g["loc"], g["attn"], nthreads=1)
    b = oracle_forward(g["value"], g["shapes"], g["loc"], g["attn"], nthreads=4)
    assert np.array_equal(a, b)
    ga = oracle_backward(g["value"], g["shapes"], g["loc"], g["attn"], g["grad_out"], nthreads=1)
    gb = oracle_backward(g["value"], g["shapes"], g["loc"], g["attn"], g["grad_out"], nthreads=4)
    for x, y in zip(ga, gb):
        assert np.array_equal(x, y)


def test_oracle_empty_queries():
    shapes = np.array([[2, 3]], dtype=np.int64)
    value = np.ones((1, 6, 1, 4), np.float32)
    loc = np.zeros((1, 0, 1, 1, 2, 2), np.float32)
    attn = np.zeros((1, 0, 1, 1, 2), np.float32)
    assert oracle_forward(value, shapes, loc, attn).shape == (1, 0, 4)
    gv, gl, ga = oracle_backward(value, shapes, loc, attn, np.zeros((1, 0, 4), np.float32))
    assert not gv.any() and gl.size == 0 and ga.size == 0


def test_layers_oracle_matches_reference_golden():
    """oracle/layers_oracle.py (CPU restatement of the reference's encoder / decoder layers) against the golden
    file produced by the reference's own classes (tests/golden/make_golden_layers.py): outputs and gradients."""
    import os
    import torch
    from oracle import layers_oracle as lo
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "layers_f32.npz"))
    t = lambda k: torch.from_numpy(z[k])
    p = {k[2:]: t(k).requires_grad_(True) for k in z.files if k.startswith("p:")}
    levels = [tuple(int(v) for v in hw) for hw in z["levels"]]
    shapes = t("levels").long()
    cfg = (8, 2, 2)                                    # heads, levels, points of the golden model
    src, pos = t("src").requires_grad_(True), t("pos").requires_grad_(True)
    tgt, qpos = t("tgt").requires_grad_(True), t("qpos").requires_grad_(True)
    memory = lo.encoder(p, "enc.", 2, src, levels, shapes, t("valid"), pos, t("mask"), cfg)
    out, ref_out = lo.decoder(p, "dec.", 2, tgt, t("ref"), memory, shapes, t("valid"), qpos, t("mask"), t("qmask"), cfg)
    _close(memory.detach().numpy(), z["memory"], 1e-4, 1e-5)
    _close(out.detach().numpy(), z["out"], 1e-4, 1e-5)
    assert np.array_equal(ref_out.numpy(), z["ref_out"])
    loss = (memory * t("w_mem")).sum() + (out * t("w_out")).sum()
    names = [k[2:] for k in z.files if k.startswith("g:")]
    grads = torch.autograd.grad(loss, [src, pos, tgt, qpos] + [p[k] for k in names])
    for k, g in zip(["g_src", "g_pos", "g_tgt", "g_qpos"] + ["g:" + k for k in names], grads):
        scale = float(np.abs(z[k]).max())
        assert float(np.abs(g.numpy() - z[k]).max()) <= 2e-4 * max(scale, 1e-6), k
