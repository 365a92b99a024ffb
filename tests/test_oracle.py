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


def _transformer_case(z, case, device="cpu"):
    import torch
    t = lambda k: torch.from_numpy(z[case + "." + k]).to(device)
    n = sum(1 for k in z.files if k.startswith(case + ".src"))
    srcs = [t("src%d" % i).requires_grad_(case == "a") for i in range(n)]
    masks, poss = [t("mask%d" % i) for i in range(n)], [t("pos%d" % i) for i in range(n)]
    sd = {k[len(case) + 3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith(case + ".p:")}
    return srcs, masks, poss, t("query_embed"), sd


def test_transformer_glue_matches_reference_golden_on_cpu():
    """kinet_b200.modules.DeformableTransformer's own logic (pyramid flattening, level embedding, valid ratios,
    per-frame encoder runs, track queries from `targets`, reference points, return tuple) against golden vectors
    from the REAL reference class (tests/golden/make_golden_transformer.py).  The encoder / decoder stacks, which
    need the CUDA kernels, are stood in for by oracle/layers_oracle.py here; tests/test_layers_gpu.py runs the
    real ones against the same file on the GPU."""
    import os
    import types
    import torch
    from kinet_b200.modules import DeformableTransformer
    from oracle import layers_oracle as lo
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "transformer_f32.npz"))

    def stand_ins(model, cfg_enc, cfg_dec):
        params = dict(model.named_parameters())

        class Enc(torch.nn.Module):
            def forward(self, src, shapes, valid, pos=None, mask=None, return_stream=False):
                levels = [tuple(hw) for hw in shapes.tolist()]
                mem = lo.encoder(params, "encoder.", model_layers[0], src, levels, shapes, valid, pos, mask, cfg_enc)
                return types.SimpleNamespace(x=mem, x_lp=mem) if return_stream else mem

        class Dec(torch.nn.Module):
            num_layers = 0

            def forward(self, tgt, ref, src, shapes, valid, query_pos=None, mask=None, qmask=None):
                memory = src.x if hasattr(src, "x") else src
                outs = []
                for i in range(model_layers[1]):
                    tgt = lo.decoder_layer(params, "decoder.layers.%d." % i, tgt, query_pos,
                                           ref[:, :, None] * valid[:, None], memory, shapes, mask, qmask, cfg_dec)
                    outs.append(tgt)
                return torch.stack(outs), torch.stack([ref] * len(outs))
        return Enc(), Dec()

    # case a: track queries, 2 levels
    srcs, masks, poss, query_embed, sd = _transformer_case(z, "a")
    model = DeformableTransformer(128, 8, 2, 2, 64, 0.0, "relu", True, 2, 2, 2)
    model.load_state_dict(sd, strict=True)
    model_layers = (2, 2)
    enc, dec = stand_ins(model, (8, 2, 2), (8, 2, 2))
    model.__dict__["_modules"]["encoder"], model.__dict__["_modules"]["decoder"] = enc, dec
    targets = [dict(track_query_hs_embeds=torch.from_numpy(z["a.track_hs"][i]), track_query_boxes=torch.from_numpy(z["a.track_boxes"][i]))
               for i in range(2)]
    hs, memory, init_ref, inter_ref, ec, eb = model(srcs, masks, poss, query_embed, targets)
    assert ec is None and eb is None
    _close(hs.detach().numpy(), z["a.hs"], 1e-4, 1e-5)
    _close(memory.detach().numpy(), z["a.memory"], 1e-4, 1e-5)
    assert np.allclose(init_ref.detach().numpy(), z["a.init_ref"], atol=1e-6)
    assert np.allclose(inter_ref.detach().numpy(), z["a.inter_ref"], atol=1e-6)

    # case b: 4 levels, separate encoder runs per frame
    srcs, masks, poss, query_embed, sd = _transformer_case(z, "b")
    model = DeformableTransformer(64, 8, 1, 1, 32, 0.0, "relu", True, 4, 2, 2, multi_frame_attention_separate_encoder=True)
    model.load_state_dict(sd, strict=True)
    model_layers = (1, 1)
    enc, dec = stand_ins(model, (8, 2, 2), (8, 4, 2))
    model.__dict__["_modules"]["encoder"], model.__dict__["_modules"]["decoder"] = enc, dec
    with torch.no_grad():
        hs, memory, init_ref, inter_ref, _, _ = model(srcs, masks, poss, query_embed, None)
    _close(hs.numpy(), z["b.hs"], 1e-4, 1e-5)
    _close(memory.numpy(), z["b.memory"], 1e-4, 1e-5)
    assert np.allclose(init_ref.numpy(), z["b.init_ref"], atol=1e-6)


def test_two_stage_transformer_glue_matches_reference_golden_on_cpu():
    """Two-stage proposals (gen_encoder_output_proposals, top-k, proposal position embedding) and the decoder's
    iterative box refinement of kinet_b200.modules against the reference class's golden case c; encoder / decoder
    layers stood in for by the oracle (the GPU twin is in tests/test_layers_gpu.py)."""
    import os
    import sys
    import types
    import torch
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from _harness import two_stage_model
    from kinet_b200.modules.deformable_transformer import inverse_sigmoid
    from oracle import layers_oracle as lo
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "transformer_f32.npz"))
    srcs, masks, poss = ([torch.from_numpy(z["c.%s%d" % (k, i)]) for i in range(2)] for k in ("src", "mask", "pos"))
    model = two_stage_model()
    params = dict(model.named_parameters())
    heads = model.decoder.bbox_embed

    class Enc(torch.nn.Module):
        def forward(self, src, shapes, valid, pos=None, mask=None, return_stream=False):
            mem = lo.encoder(params, "encoder.", 1, src, [tuple(hw) for hw in shapes.tolist()], shapes, valid, pos, mask,
                             (8, 2, 2))
            return types.SimpleNamespace(x=mem, x_lp=mem) if return_stream else mem

    class Dec(torch.nn.Module):
        num_layers = 2

        def __init__(self, real):
            super().__init__()
            self.class_embed, self.bbox_embed = real.class_embed, real.bbox_embed

        def forward(self, tgt, ref, src, shapes, valid, query_pos=None, mask=None, qmask=None):
            memory = src.x if hasattr(src, "x") else src
            outs, refs = [], []
            for i in range(2):
                rin = ref[:, :, None] * torch.cat([valid, valid], -1)[:, None]
                tgt = lo.decoder_layer(params, "decoder.layers.%d." % i, tgt, query_pos, rin, memory, shapes, mask, qmask,
                                       (8, 2, 2))
                ref = (heads[i](tgt) + inverse_sigmoid(ref)).sigmoid().detach()
                outs.append(tgt)
                refs.append(ref)
            return torch.stack(outs), torch.stack(refs)
    model.__dict__["_modules"]["encoder"], model.__dict__["_modules"]["decoder"] = Enc(), Dec(model.decoder)
    with torch.no_grad():
        hs, memory, init_ref, inter_ref, enc_cls, enc_box = model(srcs, masks, poss, None, None)
    _close(memory.numpy(), z["c.memory"], 1e-4, 1e-5)
    _close(enc_cls.numpy(), z["c.enc_cls"], 1e-4, 1e-5)
    assert np.array_equal(np.isinf(enc_box.numpy()), np.isinf(z["c.enc_box"]))
    fin = np.isfinite(z["c.enc_box"])
    assert np.allclose(enc_box.numpy()[fin], z["c.enc_box"][fin], rtol=1e-4, atol=1e-4)
    assert np.allclose(init_ref.numpy(), z["c.init_ref"], atol=1e-5)
    _close(hs.numpy(), z["c.hs"], 1e-4, 1e-5)
    assert np.allclose(inter_ref.numpy(), z["c.inter_ref"], atol=1e-5)
