"""GPU parity of the layer glue kernels (include/kinet_layers_b200.h) and of the encoder / decoder classes
(kinet_b200/modules/deformable_transformer.py) against (a) fp64 PyTorch formulas of the same arithmetic and
(b) golden vectors produced by the REFERENCE's own classes (tests/golden/make_golden_layers.py)."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _dev():
    return torch.device("cuda", 0)


def _rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("C", [4, 36, 128, 256, 288, 512, 1024])
@pytest.mark.parametrize("stream,lp", [(torch.float32, torch.float32), (torch.float32, torch.bfloat16),
                                       (torch.bfloat16, torch.bfloat16), (torch.float32, torch.float16)])
@pytest.mark.parametrize("pos_mode", ["none", "full", "batch_broadcast"])
def test_add_layer_norm_forward_backward(C, stream, lp, pos_mode):
    from kinet_b200.functions.layer_ops import AddLayerNormFunction
    dev = _dev()
    g = torch.Generator(device=dev).manual_seed(C)
    N, Lq = 3, 13
    x = torch.randn(N, Lq, C, device=dev, generator=g).to(stream).requires_grad_(True)
    r = (0.5 * torch.randn(N, Lq, C, device=dev, generator=g)).to(lp).requires_grad_(True)
    gamma = (1 + 0.2 * torch.randn(C, device=dev, generator=g)).requires_grad_(True)
    beta = (0.2 * torch.randn(C, device=dev, generator=g)).requires_grad_(True)
    pos = None
    if pos_mode == "full":
        pos = (0.3 * torch.randn(N, Lq, C, device=dev, generator=g)).requires_grad_(True)
    elif pos_mode == "batch_broadcast":
        pos = (0.3 * torch.randn(Lq, C, device=dev, generator=g)).requires_grad_(True)
    pos_in = pos if pos_mode != "batch_broadcast" else pos[None].expand(N, -1, -1)
    want_lp = lp != stream
    y, y_lp, ypos = AddLayerNormFunction.apply(x, r, pos_in, gamma, beta, 1e-5, lp, want_lp)
    assert y.dtype == stream and (y_lp is None) == (not want_lp) and (ypos is None) == (pos is None)

    # fp64 restatement on the same (rounded) inputs
    xd, rd = x.detach().double().requires_grad_(True), r.detach().double().requires_grad_(True)
    gd, bd = gamma.detach().double().requires_grad_(True), beta.detach().double().requires_grad_(True)
    pd = pos.detach().double().requires_grad_(True) if pos is not None else None
    yd = F.layer_norm(xd + rd, (C,), gd, bd, 1e-5)
    tol_s = 1e-5 if stream == torch.float32 else 1.2e-2
    tol_l = 1e-5 if lp == torch.float32 else 1.2e-2
    assert _rel(y, yd) < tol_s
    ws = [torch.randn(N, Lq, C, device=dev, generator=g) for _ in range(3)]
    loss = (y.float() * ws[0]).sum()
    lossd = (yd * ws[0].double()).sum()
    if want_lp:
        assert _rel(y_lp, yd) < tol_l
        loss = loss + (y_lp.float() * ws[1]).sum()
        lossd = lossd + (yd * ws[1].to(lp).double()).sum()     # the gradient reaches the kernel rounded to lp
    if pos is not None:
        assert _rel(ypos, yd + pd) < tol_l
        loss = loss + (ypos.float() * ws[2]).sum()
        lossd = lossd + ((yd + pd) * ws[2].to(lp).double()).sum()
    loss.backward()
    lossd.backward()
    gt = 2e-5 if (stream == torch.float32 and lp == torch.float32) else 1.5e-2
    assert _rel(x.grad, xd.grad) < (2e-5 if stream == torch.float32 and lp == torch.float32 else 1.5e-2)
    assert _rel(r.grad, rd.grad) < gt
    assert _rel(gamma.grad, gd.grad) < gt and _rel(beta.grad, bd.grad) < gt
    if pos is not None:
        assert pos.grad.shape == pos.shape and _rel(pos.grad, pd.grad) < gt


def test_add_layer_norm_no_branch_many_rows_and_determinism():
    from kinet_b200.functions.layer_ops import AddLayerNormFunction
    dev = _dev()
    g = torch.Generator(device=dev).manual_seed(1)
    x = torch.randn(2, 22223, 256, device=dev, generator=g).requires_grad_(True)
    gamma = torch.randn(256, device=dev, generator=g).requires_grad_(True)
    beta = torch.randn(256, device=dev, generator=g).requires_grad_(True)
    w = torch.randn(2, 22223, 256, device=dev, generator=g)
    grads = []
    for _ in range(2):
        y, _, _ = AddLayerNormFunction.apply(x, None, None, gamma, beta, 1e-5, torch.float32, False)
        grads.append(torch.autograd.grad((y * w).sum(), (x, gamma, beta)))
    ref = F.layer_norm(x.double(), (256,), gamma.double(), beta.double(), 1e-5)
    assert _rel(y, ref) < 1e-5
    rg = torch.autograd.grad((ref * w.double()).sum(), (x, gamma, beta))
    for a, b in zip(grads[0], rg):
        assert _rel(a, b) < 2e-5
    for a, b in zip(grads[0], grads[1]):          # fixed summation order: bit-identical from run to run
        assert torch.equal(a, b)


@pytest.mark.parametrize("rows,C", [(1, 4), (77, 128), (600, 256), (2048, 288), (2049, 256), (44446, 256), (44446, 1024),
                                     (20906, 288), (5, 1000)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
def test_column_sum(rows, C, dtype):
    from kinet_b200.functions.layer_ops import column_sum
    dev = _dev()
    a = torch.randn(rows, C, device=dev, generator=torch.Generator(device=dev).manual_seed(rows)).to(dtype)
    out = column_sum(a, torch.float32)
    ref = a.double().sum(0)
    assert float((out.double() - ref).abs().max()) <= 1e-5 * float(a.double().abs().sum(0).max())
    assert torch.equal(out, column_sum(a, torch.float32))
    assert column_sum(a).dtype == dtype


@pytest.mark.parametrize("rows", [3, 600, 2048, 2049, 9000])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_relu_backward_column_sum(rows, dtype):
    from kinet_b200.functions.layer_ops import relu_backward_column_sum
    dev = _dev()
    g = torch.Generator(device=dev).manual_seed(rows)
    dy = torch.randn(rows, 1024, device=dev, generator=g).to(dtype)
    h = torch.relu(torch.randn(rows, 1024, device=dev, generator=g)).to(dtype)
    dh, db = relu_backward_column_sum(dy, h, torch.float32)
    ref = dy * (h > 0).to(dtype)
    assert torch.equal(dh, ref)
    assert float((db.double() - ref.double().sum(0)).abs().max()) <= 1e-5 * float(ref.double().abs().sum(0).max())


@pytest.mark.parametrize("amp", [False, True])
def test_linear_function_matches_nn_linear(amp):
    from kinet_b200.functions.layer_ops import linear
    dev = _dev()
    torch.manual_seed(0)
    lin = torch.nn.Linear(256, 384).to(dev)
    x = torch.randn(2, 777, 256, device=dev, requires_grad=True)
    w = torch.randn(2, 777, 384, device=dev)
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
        y = linear(x, lin)
        y0 = lin(x)
    assert y.dtype == y0.dtype and _rel(y, y0) < (1e-2 if amp else 1e-5)
    g = torch.autograd.grad((y.float() * w).sum(), (x, lin.weight, lin.bias))
    g0 = torch.autograd.grad((y0.float() * w).sum(), (x, lin.weight, lin.bias))
    tol = 2e-2 if amp else 1e-5
    for a, b in zip(g, g0):
        assert a.dtype == b.dtype and _rel(a, b) < tol
    # the bias gradient is an fp32 sum of the (rounded) upstream gradient: compare with fp64
    up = w.to(y.dtype).double().sum((0, 1))
    assert _rel(g[2], up) < 1e-5


@pytest.mark.parametrize("amp", [False, True])
def test_linear_relu_fused(amp):
    """ReLU in the GEMM epilogue forward, ReLU backward fused with the bias-gradient column sum."""
    from kinet_b200.functions.layer_ops import linear
    dev = _dev()
    torch.manual_seed(1)
    lin = torch.nn.Linear(256, 1024).to(dev)
    x = torch.randn(2, 555, 256, device=dev, requires_grad=True)
    w = torch.randn(2, 555, 1024, device=dev)
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
        y = linear(x, lin, relu=True)
        y0 = torch.relu(lin(x))
        with torch.no_grad():
            yi = linear(x, lin, relu=True)          # inference path: cached low-precision weights + epilogue ReLU
    tol = 2e-2 if amp else 1e-5
    assert y.dtype == y0.dtype == yi.dtype and _rel(y, y0) < tol and _rel(yi, y0) < tol
    assert float(y.min()) == 0.0
    g = torch.autograd.grad((y.float() * w).sum(), (x, lin.weight, lin.bias))
    g0 = torch.autograd.grad((y0.float() * w).sum(), (x, lin.weight, lin.bias))
    for a, b in zip(g, g0):
        assert a.dtype == b.dtype and _rel(a, b) < (3e-2 if amp else 1e-5)


@pytest.mark.parametrize("masked", [False, True])
def test_self_attention_matches_nn_multihead_attention(masked):
    from kinet_b200.modules.deformable_transformer import _self_attention
    dev = _dev()
    torch.manual_seed(2)
    mha = torch.nn.MultiheadAttention(128, 8, dropout=0.0).to(dev)
    with torch.no_grad():
        mha.in_proj_bias.normal_(0, 0.1)
        mha.out_proj.bias.normal_(0, 0.1)
    q = torch.randn(2, 37, 128, device=dev, requires_grad=True)
    v = torch.randn(2, 37, 128, device=dev, requires_grad=True)
    mask = None
    if masked:
        mask = torch.zeros(2, 37, dtype=torch.bool, device=dev)
        mask[0, -5:] = True
    out = _self_attention(mha, q, v, mask)
    ref = mha(q.transpose(0, 1), q.transpose(0, 1), v.transpose(0, 1), key_padding_mask=mask)[0].transpose(0, 1)
    assert _rel(out, ref) < 1e-5
    w = torch.randn_like(out)
    ps = (q, v, mha.in_proj_weight, mha.in_proj_bias, mha.out_proj.weight, mha.out_proj.bias)
    g = torch.autograd.grad((out * w).sum(), ps)
    g0 = torch.autograd.grad((ref * w).sum(), ps)
    for a, b in zip(g, g0):
        assert _rel(a, b) < 2e-5


@pytest.mark.parametrize("heads,C", [(8, 288), (8, 256)])
def test_self_attention_mask_rides_in_a_spare_head_channel_bf16(heads, C):
    """16-bit masked self-attention (key-padding mask folded into a spare head channel so that SDPA stays on the
    flash kernel; head_dim 36 and 32) against nn.MultiheadAttention under the same autocast, incl. gradients."""
    from kinet_b200.modules.deformable_transformer import _self_attention
    dev = _dev()
    torch.manual_seed(3)
    mha = torch.nn.MultiheadAttention(C, heads, dropout=0.0).to(dev)
    with torch.no_grad():
        mha.in_proj_bias.normal_(0, 0.1)
    q = torch.randn(2, 70, C, device=dev, requires_grad=True)
    v = torch.randn(2, 70, C, device=dev, requires_grad=True)
    mask = torch.zeros(2, 70, dtype=torch.bool, device=dev)
    mask[0, -9:] = True
    mask[1, 3] = True
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out = _self_attention(mha, q.bfloat16(), v.bfloat16(), mask)
        ref = mha(q.transpose(0, 1), q.transpose(0, 1), v.transpose(0, 1), key_padding_mask=mask)[0].transpose(0, 1)
    assert out.dtype == torch.bfloat16 and bool(torch.isfinite(out).all())
    assert _rel(out, ref) < 3e-2
    # masked keys get exactly zero weight: changing their content does not change the output
    v2 = v.detach().clone()
    v2[0, -9:] += 100.0
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        a = _self_attention(mha, q.detach().bfloat16(), v.detach().bfloat16(), mask)
        b = _self_attention(mha, q.detach().bfloat16(), v2.bfloat16(), mask)
    assert torch.equal(a, b)
    w = torch.randn_like(out, dtype=torch.float32)
    g = torch.autograd.grad((out.float() * w).sum(), (q, v, mha.in_proj_weight))
    g0 = torch.autograd.grad((ref.float() * w).sum(), (q, v, mha.in_proj_weight))
    for x, y in zip(g, g0):
        assert _rel(x, y) < 5e-2


def test_flat_params_training_step_matches_plain_parameters():
    """FlatParams (flat master / gradient buffers + bf16 mirror) gives the gradients and the AdamW update of the
    same model with ordinary parameters."""
    from kinet_b200.training import FlatParams
    z = _load_golden()
    dev = _dev()
    i = _golden_inputs(z, dev)

    def one_step(flat):
        enc, _ = _build_from_golden(z, dev)
        enc.train()
        if flat:
            fp = FlatParams(enc)
            opt = torch.optim.AdamW([fp.flat_parameter], lr=1e-3, weight_decay=1e-2)
        else:
            opt = torch.optim.AdamW(enc.parameters(), lr=1e-3, weight_decay=1e-2)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            mem = enc(i["src"].detach(), i["shapes"], i["valid"], i["pos"].detach(), i["mask"])
        (mem.float() * i["w_mem"]).sum().backward()
        grads = {k: p.grad.clone() for k, p in enc.named_parameters()}
        opt.step()
        if flat:
            fp.refresh()
            lin = enc.layers[0].linear1
            assert torch.equal(lin._kinet_mirror[0], lin.weight.detach().bfloat16())
        return grads, {k: p.detach().clone() for k, p in enc.named_parameters()}
    ga, pa = one_step(True)
    gb, pb = one_step(False)
    for k in gb:
        assert _rel(ga[k], gb[k]) < 3e-2, k
        assert _rel(pa[k], pb[k]) < 1e-3, k


def test_graphed_train_step_follows_eager_training():
    """training.GraphedTrainStep (two CUDA graphs around the gradient exchange) moves the parameters like the
    same steps run eagerly with ordinary parameters."""
    from kinet_b200.training import GraphedTrainStep
    z = _load_golden()
    dev = _dev()
    i = _golden_inputs(z, dev)
    batches = [(torch.from_numpy(z["src"]).to(dev) * (1 + 0.1 * k), torch.from_numpy(z["pos"]).to(dev)) for k in range(3)]

    def make():
        enc, _ = _build_from_golden(z, dev)
        enc.train()

        def loss_fn(src, pos):
            with torch.autocast("cuda", dtype=torch.bfloat16):
                return (enc(src, i["shapes"], i["valid"], pos, i["mask"]).float() * i["w_mem"]).sum() * 1e-3
        return enc, loss_fn
    enc_a, loss_a = make()
    opt = torch.optim.AdamW(enc_a.parameters(), lr=1e-3, weight_decay=1e-2)
    losses_a = []
    for k in range(3 + 3):                      # GraphedTrainStep warms up with 3 real steps on its example batch
        src, pos = batches[0] if k < 3 else batches[k - 3]
        opt.zero_grad()
        loss = loss_a(src, pos)
        loss.backward()
        opt.step()
        losses_a.append(float(loss))
    enc_b, loss_b = make()
    step = GraphedTrainStep(enc_b, loss_b, lambda p: torch.optim.AdamW(p, lr=1e-3, weight_decay=1e-2, capturable=True),
                            batches[0])
    losses_b = [float(step(*batches[k])) for k in range(3)]
    for a, b in zip(losses_a[3:], losses_b):
        assert abs(a - b) <= 2e-2 * max(abs(a), 1e-3), (losses_a, losses_b)
    for (k, pa), (_, pb) in zip(enc_a.named_parameters(), enc_b.named_parameters()):
        assert _rel(pb.detach(), pa.detach()) < 2e-2, k


def _load_golden():
    z = np.load(os.path.join(HERE, "golden", "layers_f32.npz"))
    return {k: z[k] for k in z.files}


def _build_from_golden(z, dev):
    from kinet_b200.modules import (DeformableTransformerDecoder, DeformableTransformerDecoderLayer,
                                    DeformableTransformerEncoder, DeformableTransformerEncoderLayer)
    d, ffn, L, M, P = 128, 64, 2, 8, 2
    enc = DeformableTransformerEncoder(DeformableTransformerEncoderLayer(d, ffn, 0.0, "relu", L, M, P), 2)
    dec = DeformableTransformerDecoder(DeformableTransformerDecoderLayer(d, ffn, 0.0, "relu", L, M, P), 2)
    for prefix, m in (("enc.", enc), ("dec.", dec)):
        sd = {k[len("p:" + prefix):]: torch.from_numpy(v) for k, v in z.items() if k.startswith("p:" + prefix)}
        missing, unexpected = m.load_state_dict(sd, strict=True)
        assert not missing and not unexpected
    return enc.to(dev).eval(), dec.to(dev).eval()


def _golden_inputs(z, dev):
    t = lambda k: torch.from_numpy(z[k]).to(dev)
    return dict(shapes=t("levels").long(), src=t("src").requires_grad_(True), pos=t("pos").requires_grad_(True),
                valid=t("valid"), mask=t("mask"), tgt=t("tgt").requires_grad_(True),
                qpos=t("qpos").requires_grad_(True), ref=t("ref"), qmask=t("qmask"), w_mem=t("w_mem"), w_out=t("w_out"))


def test_encoder_decoder_match_reference_golden_fp32():
    """The reference's own classes (CPU, fp32, oracle op) produced the golden file; ours run on the GPU."""
    z = _load_golden()
    dev = _dev()
    enc, dec = _build_from_golden(z, dev)
    i = _golden_inputs(z, dev)
    memory = enc(i["src"], i["shapes"], i["valid"], i["pos"], i["mask"])
    out, ref_out = dec(i["tgt"], i["ref"], memory, i["shapes"], i["valid"], i["qpos"], i["mask"], i["qmask"])
    assert _rel(memory, torch.from_numpy(z["memory"]).to(dev)) < 2e-5
    assert _rel(out, torch.from_numpy(z["out"]).to(dev)) < 2e-5
    assert torch.equal(ref_out.cpu(), torch.from_numpy(z["ref_out"]))
    loss = (memory * i["w_mem"]).sum() + (out * i["w_out"]).sum()
    params = dict(("enc." + k, v) for k, v in enc.named_parameters())
    params.update(("dec." + k, v) for k, v in dec.named_parameters())
    names = [k for k in params if ("g:" + k) in z]
    assert len(names) >= 60
    grads = torch.autograd.grad(loss, [i["src"], i["pos"], i["tgt"], i["qpos"]] + [params[k] for k in names])
    for k, g in zip(["g_src", "g_pos", "g_tgt", "g_qpos"] + ["g:" + k for k in names], grads):
        assert _rel(g, torch.from_numpy(z[k]).to(dev)) < 1e-4, k


def test_single_layer_forward_signatures_match_stream_path():
    """layer.forward(...) with the reference's signature equals the stacked (stream) execution."""
    z = _load_golden()
    dev = _dev()
    enc, dec = _build_from_golden(z, dev)
    i = _golden_inputs(z, dev)
    with torch.no_grad():
        refp = enc.get_reference_points(i["shapes"], i["valid"], dev)
        x = i["src"]
        for layer in enc.layers:
            x = layer(x, i["pos"], refp, i["shapes"], i["mask"])
        memory = enc(i["src"], i["shapes"], i["valid"], i["pos"], i["mask"])
        assert _rel(x, memory) < 1e-6
        st = enc(i["src"], i["shapes"], i["valid"], i["pos"], i["mask"], return_stream=True)
        out_a, _ = dec(i["tgt"], i["ref"], st, i["shapes"], i["valid"], i["qpos"], i["mask"], i["qmask"])
        t = i["tgt"]
        rin = i["ref"][:, :, None] * i["valid"][:, None]
        for layer in dec.layers:
            t = layer(t, i["qpos"], rin, memory, i["shapes"], i["mask"], i["qmask"])
        assert _rel(t, out_a) < 1e-6
        f = enc.layers[0].forward_ffn(i["src"])
        l0 = enc.layers[0]
        f0 = l0.norm2(i["src"] + l0.linear2(torch.relu(l0.linear1(i["src"]))))
        assert _rel(f, f0) < 1e-5


def test_autocast_matches_unfused_composition():
    """Under bf16 autocast the fused layers follow what the reference's modules would compute under autocast
    (fp32 residual stream, bf16 GEMM operands): compare with that composition written in plain PyTorch."""
    z = _load_golden()
    dev = _dev()
    enc, _ = _build_from_golden(z, dev)
    i = _golden_inputs(z, dev)
    layer = enc.layers[0]
    refp = enc.get_reference_points(i["shapes"], i["valid"], dev)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        y = layer(i["src"], i["pos"], refp, i["shapes"], i["mask"])
        src = i["src"]
        src2 = layer.self_attn(src + i["pos"], refp, src, i["shapes"], i["mask"])
        s = layer.norm1(src + src2)
        y0 = layer.norm2(s + layer.linear2(torch.relu(layer.linear1(s))))
    assert y.dtype == torch.float32 and y0.dtype == torch.float32
    assert _rel(y, y0) < 2e-2
    g = torch.autograd.grad(y.square().sum(), [i["src"], layer.norm2.weight, layer.linear1.bias], retain_graph=True)
    g0 = torch.autograd.grad(y0.square().sum(), [i["src"], layer.norm2.weight, layer.linear1.bias])
    for a, b in zip(g, g0):
        assert _rel(a, b) < 5e-2


def test_cpu_tensors_raise():
    from kinet_b200.functions.layer_ops import AddLayerNormFunction, column_sum
    with pytest.raises(RuntimeError, match="Not implemented on the CPU"):
        AddLayerNormFunction.apply(torch.zeros(2, 8), None, None, torch.ones(8), torch.zeros(8), 1e-5, torch.float32, False)
    with pytest.raises(RuntimeError, match="Not implemented on the CPU"):
        column_sum(torch.zeros(2, 8))


def test_layer_abi_from_a_cpp_host_program(tmp_path):
    """include/kinet_layers_b200.h is a C ABI: tests/abi_c/layers_smoke.cpp links libmsda_b200.so and the CUDA
    runtime only (no torch, no Python in the process), runs add+LayerNorm forward / backward and a column sum and
    checks them against its own double-precision restatement."""
    import shutil
    import subprocess
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available on this box")
    root = os.path.dirname(HERE)
    libdir = os.path.join(root, "kinet_b200")
    exe = str(tmp_path / "layers_smoke")
    env = {k: v for k, v in os.environ.items() if k not in ("CC", "CXX")}
    subprocess.check_call([nvcc, "-O2", "-std=c++17", "-Wno-deprecated-gpu-targets", "-I", os.path.join(root, "include"),
                           "-o", exe, os.path.join(root, "tests", "abi_c", "layers_smoke.cpp"),
                           "-L", libdir, "-lmsda_b200", "-Xlinker", "-rpath", "-Xlinker", libdir], env=env)
    res = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    print(res.stdout)
    assert res.returncode == 0, res.stdout + res.stderr
    assert "PASS" in res.stdout


def _transformer_golden(case, dev):
    z = np.load(os.path.join(HERE, "golden", "transformer_f32.npz"))
    t = lambda k: torch.from_numpy(z[case + "." + k]).to(dev)
    n = sum(1 for k in z.files if k.startswith(case + ".src"))
    srcs = [t("src%d" % i).requires_grad_(case == "a") for i in range(n)]
    masks, poss = [t("mask%d" % i) for i in range(n)], [t("pos%d" % i) for i in range(n)]
    sd = {k[len(case) + 3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith(case + ".p:")}
    return z, t, srcs, masks, poss, sd


def test_deformable_transformer_matches_reference_golden_track_queries():
    """kinet_b200.modules.DeformableTransformer on the GPU against the REAL reference DeformableTransformer
    (CPU, oracle op; tests/golden/make_golden_transformer.py): 2 levels, 7 object + 3 track queries from `targets`,
    intermediate outputs, input gradients."""
    from kinet_b200.modules import DeformableTransformer
    dev = _dev()
    z, t, srcs, masks, poss, sd = _transformer_golden("a", dev)
    model = DeformableTransformer(128, 8, 2, 2, 64, 0.0, "relu", True, 2, 2, 2)
    model.load_state_dict(sd, strict=True)
    model = model.to(dev).eval()
    query_embed = t("query_embed").requires_grad_(True)
    targets = [dict(track_query_hs_embeds=t("track_hs")[i], track_query_boxes=t("track_boxes")[i]) for i in range(2)]
    hs, memory, init_ref, inter_ref, ec, eb = model(srcs, masks, poss, query_embed, targets)
    assert ec is None and eb is None and hs.shape == (2, 2, 10, 128)
    assert _rel(hs, t("hs")) < 2e-5 and _rel(memory, t("memory")) < 2e-5
    assert _rel(init_ref, t("init_ref")) < 1e-6 and _rel(inter_ref, t("inter_ref")) < 1e-6
    grads = torch.autograd.grad((hs * t("w_hs")).sum() + (memory * t("w_mem")).sum(), srcs + [query_embed])
    for g, k in zip(grads, ["g_src0", "g_src1", "g_query_embed"]):
        assert _rel(g, t(k)) < 1e-4, k
    # the level-shape tensor is reused from call to call (per-tensor caches hit; CUDA-graph capturable)
    from kinet_b200.modules.deformable_transformer import _shapes_tensor
    assert _shapes_tensor([(6, 8), (3, 4)], srcs[0].device) is _shapes_tensor([(6, 8), (3, 4)], srcs[0].device)


def test_deformable_transformer_matches_reference_golden_separate_frame_encoders():
    """4 levels (2 per frame), multi_frame_attention_separate_encoder=True: the encoder runs once per frame."""
    from kinet_b200.modules import DeformableTransformer
    dev = _dev()
    z, t, srcs, masks, poss, sd = _transformer_golden("b", dev)
    model = DeformableTransformer(64, 8, 1, 1, 32, 0.0, "relu", True, 4, 2, 2, multi_frame_attention_separate_encoder=True)
    model.load_state_dict(sd, strict=True)
    model = model.to(dev).eval()
    with torch.no_grad():
        hs, memory, init_ref, inter_ref, _, _ = model(srcs, masks, poss, t("query_embed"), None)
    assert _rel(hs, t("hs")) < 2e-5 and _rel(memory, t("memory")) < 2e-5
    assert _rel(init_ref, t("init_ref")) < 1e-6


def test_deformable_transformer_two_stage_matches_reference_golden():
    """Two-stage proposals + iterative box refinement (4-d reference points through the fused MSDeformAttn
    prologue) on the GPU against the reference class's golden case c."""
    import sys
    sys.path.insert(0, HERE)
    from _harness import two_stage_model
    dev = _dev()
    z = np.load(os.path.join(HERE, "golden", "transformer_f32.npz"))
    srcs, masks, poss = ([torch.from_numpy(z["c.%s%d" % (k, i)]).to(dev) for i in range(2)] for k in ("src", "mask", "pos"))
    model = two_stage_model().to(dev)
    with torch.no_grad():
        hs, memory, init_ref, inter_ref, enc_cls, enc_box = model(srcs, masks, poss, None, None)
    t = lambda k: torch.from_numpy(z["c." + k]).to(dev)
    assert _rel(memory, t("memory")) < 2e-5 and _rel(enc_cls, t("enc_cls")) < 2e-5
    assert torch.equal(torch.isinf(enc_box), torch.isinf(t("enc_box")))
    # the same proposals were selected (top-k on the class logit) and refined to the same boxes
    assert _rel(init_ref, t("init_ref")) < 1e-5
    assert _rel(hs, t("hs")) < 5e-5 and _rel(inter_ref, t("inter_ref")) < 1e-5


@pytest.mark.gpu
@pytest.mark.parametrize("C,stream,lp", [(256, torch.float32, torch.bfloat16), (288, torch.float32, torch.float32),
                                         (1024, torch.bfloat16, torch.bfloat16)])
def test_add_layer_norm_fused_dropout_and_branch_bias(C, stream, lp):
    """y = LayerNorm(x + dropout(r)) with the mask generated inside the kernels: the mask the backward regenerates
    (read off d r) is the one the forward used, its density is p, the arithmetic matches PyTorch's with that mask, the
    column sums of d r come out as the branch bias gradient, new sites / steps draw new masks."""
    from kinet_b200.functions import layer_ops as lo
    dev = _dev()
    g = torch.Generator().manual_seed(C)
    rows, p = 3000, 0.1
    x = torch.randn(rows, C, generator=g).to(dev, stream).requires_grad_(True)
    r = torch.randn(rows, C, generator=g).to(dev, lp).requires_grad_(True)
    gamma = (1 + 0.1 * torch.randn(C, generator=g)).to(dev).requires_grad_(True)
    beta = (0.1 * torch.randn(C, generator=g)).to(dev).requires_grad_(True)
    bias = torch.zeros(C, device=dev, requires_grad=True)
    w = torch.randn(rows, C, generator=g).to(dev)
    lo.set_dropout_seed(1234)
    y, _, _ = lo.AddLayerNormFunction.apply(x, r, None, gamma, beta, 1e-5, lp, False, p, bias)
    dx, dr, dg, db, dbias = torch.autograd.grad((y.float() * w).sum(), [x, r, gamma, beta, bias])
    # the mask, as the backward saw it: d r = mask / (1 - p') * d x
    mask = dr.float() != 0
    dropped = 1.0 - mask.float().mean().item()
    assert abs(dropped - p) < 4 * (p * (1 - p) / (rows * C)) ** 0.5 + 2e-5, dropped
    scale = 65536.0 / (65536 - round(p * 65536))
    xd, rd, gd, bd = (t.detach().double().requires_grad_(True) for t in (x, r, gamma, beta))
    z = xd + rd * mask.double() * scale
    yd = torch.nn.functional.layer_norm(z, (C,), gd, bd, 1e-5)
    rdx, rdr, rdg, rdb = torch.autograd.grad((yd * w.double()).sum(), [xd, rd, gd, bd])
    tol = 2e-5 if stream == torch.float32 else 1.2e-2
    assert _rel(y, yd.detach()) < tol                 # the forward used the same mask
    assert _rel(dx, rdx) < tol and _rel(dg, rdg) < max(tol, 1e-4) and _rel(db, rdb) < max(tol, 1e-4)
    assert _rel(dr, rdr) < (2e-5 if lp == torch.float32 else 1.2e-2)
    assert _rel(dbias, rdr.sum(0)) < (1e-4 if lp == torch.float32 else 1.2e-2)
    # a second call is a new site: another mask; the same site and step reproduce (checked through the C ABI's determinism:
    # forward twice with the site counter rewound)
    site = lo._drop_state["site"]
    y2, _, _ = lo.AddLayerNormFunction.apply(x, r, None, gamma, beta, 1e-5, lp, False, p, None)
    assert not torch.equal(y2, y)
    lo._drop_state["site"] = site - 1
    y3, _, _ = lo.AddLayerNormFunction.apply(x, r, None, gamma, beta, 1e-5, lp, False, p, None)
    assert torch.equal(y3, y)
    lo._drop_state["site"] = site - 1
    lo.advance_dropout_step(dev)
    y4, _, _ = lo.AddLayerNormFunction.apply(x, r, None, gamma, beta, 1e-5, lp, False, p, None)
    assert not torch.equal(y4, y)
    # p = 0: the plain kernel, bit for bit
    y0, _, _ = lo.AddLayerNormFunction.apply(x, r, None, gamma, beta, 1e-5, lp, False, 0.0, bias)
    y1, _, _ = lo.AddLayerNormFunction.apply(x, r, None, gamma, beta, 1e-5, lp, False)
    assert torch.equal(y0, y1)
    d0 = torch.autograd.grad((y0.float() * w).sum(), [x, r, bias])
    d1 = torch.autograd.grad((y1.float() * w).sum(), [x, r])
    assert torch.equal(d0[0], d1[0]) and torch.equal(d0[1], d1[1])
    assert _rel(d0[2], d1[1].float().sum(0)) < (1e-5 if lp == torch.float32 else 4e-3)


@pytest.mark.gpu
def test_layers_bias_gradients_come_out_of_the_norm_steps():
    """Encoder / decoder layers route the bias gradients of output_proj / out_proj / linear2 through the fused add +
    LayerNorm backward (no column-sum pass of their own): every parameter still gets its gradient, equal to the plain
    composition's (dropout 0), and training mode with dropout 0.1 runs with finite, non-zero gradients everywhere."""
    from kinet_b200.modules import DeformableTransformerDecoderLayer, DeformableTransformerEncoderLayer
    dev = _dev()
    torch.manual_seed(3)
    d, L = 128, 2
    levels = [(6, 8), (3, 4)]
    S = sum(h * w for h, w in levels)
    shapes = torch.as_tensor(levels, device=dev)
    for p_drop in (0.0, 0.1):
        enc = DeformableTransformerEncoderLayer(d, 48, p_drop, "relu", L, 8, 2).to(dev).train()
        dec = DeformableTransformerDecoderLayer(d, 48, p_drop, "relu", L, 8, 2).to(dev).train()
        with torch.no_grad():
            for m in (enc, dec):
                for n_, p_ in m.named_parameters():
                    if p_.dim() == 1 and "norm" not in n_:
                        p_.normal_(0, 0.1)
        src = torch.randn(2, S, d, device=dev, requires_grad=True)
        pos = torch.randn(2, S, d, device=dev)
        ref_e = torch.rand(2, S, L, 2, device=dev)
        tgt = torch.randn(2, 5, d, device=dev)
        qpos = torch.randn(2, 5, d, device=dev)
        ref_d = torch.rand(2, 5, L, 2, device=dev)
        mem = enc(src, pos, ref_e, shapes)
        out = dec(tgt, qpos, ref_d, mem, shapes)
        out.square().sum().backward()
        for m in (enc, dec):
            for n_, p_ in m.named_parameters():
                assert p_.grad is not None and torch.isfinite(p_.grad).all(), n_
                if "sampling_offsets" not in n_ and "attention_weights" not in n_:
                    assert p_.grad.abs().max() > 0, n_
        if p_drop == 0.0:
            # the same layers with every bias gradient computed by its own Linear layer (skip flags off)
            import kinet_b200.functions.layer_ops as lo
            got = {n_: p_.grad.clone() for m, tag in ((enc, "e."), (dec, "d.")) for n_, p_ in
                   ((tag + k, v) for k, v in m.named_parameters())}
            for m in (enc, dec):
                m.zero_grad()
            src.grad = None
            orig = lo.LinearFunction.forward
            real_apply = lo.AddLayerNormFunction.apply
            lo_linear_params = lo.linear_params
            try:
                lo.linear_params = lambda x, w, b, relu=False, mirror=None, skip_bias_grad=False, drop_p=0.0: lo_linear_params(x, w, b, relu, mirror, False, drop_p)
                import kinet_b200.modules.deformable_transformer as dt
                dt_norm = dt.add_layer_norm
                dt.add_layer_norm = lambda x, r, norm, pos=None, lp_dtype=None, want_lp=False, drop_p=0.0, branch_bias=None: \
                    dt_norm(x, r, norm, pos, lp_dtype, want_lp, drop_p, None)
                lo_linear = lo.linear
                mem = enc(src, pos, ref_e, shapes)
                out = dec(tgt, qpos, ref_d, mem, shapes)
                out.square().sum().backward()
            finally:
                lo.linear_params = lo_linear_params
                dt.add_layer_norm = dt_norm
            for m, tag in ((enc, "e."), (dec, "d.")):
                for n_, p_ in m.named_parameters():
                    assert _rel(got[tag + n_], p_.grad) < 1e-4, n_


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_linear_relu_dropout_without_a_mask_tensor(dtype):
    """y = dropout(relu(x W^T + b)) through LinearFunction: density of the zeros, and the backward -- which only sees
    y's sign pattern -- against autograd through the same mask."""
    from kinet_b200.functions import layer_ops as lo
    dev = _dev()
    g = torch.Generator().manual_seed(2)
    rows, cin, cout, p = 2500, 64, 256, 0.1
    lin = torch.nn.Linear(cin, cout).to(dev)
    x = torch.randn(rows, cin, generator=g).to(dev).requires_grad_(True)
    w = torch.randn(rows, cout, generator=g).to(dev)
    with torch.autocast("cuda", dtype=dtype, enabled=dtype != torch.float32):
        y = lo.linear(x, lin, relu=True, drop_p=p)
    gx, gw, gb = torch.autograd.grad((y.float() * w).sum(), [x, lin.weight, lin.bias])
    with torch.no_grad(), torch.autocast("cuda", dtype=dtype, enabled=dtype != torch.float32):
        h = torch.relu(lin(x))
    open_ = h > 0
    kept = (y != 0) & open_
    frac = 1.0 - kept.sum().item() / open_.sum().item()
    assert abs(frac - p) < 5 * (p * (1 - p) / open_.sum().item()) ** 0.5 + 1e-4, frac
    scale = lo.dropout_scale(p)
    tol = 2e-5 if dtype == torch.float32 else 1.2e-2
    assert _rel(y.float(), (h.float() * kept * scale)) < tol
    xd = x.detach().double().requires_grad_(True)
    wd, bd = lin.weight.detach().double().requires_grad_(True), lin.bias.detach().double().requires_grad_(True)
    # (the surviving AND open elements, as y itself shows them: in bf16 the ReLU of a pre-activation within rounding
    # distance of 0 may open in one arithmetic and not in the other, which is not what this test is about)
    yd = torch.nn.functional.linear(xd, wd, bd) * (y != 0).double() * scale
    rx, rw, rb = torch.autograd.grad((yd * w.double()).sum(), [xd, wd, bd])
    assert _rel(gx, rx) < tol and _rel(gw, rw) < tol and _rel(gb, rb) < tol
