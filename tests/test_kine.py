"""The literal KineT (`kine`) path -- SURVEY.md section 8 row f4: kinet_b200.modules.kinematic_transformer and the
short-sequence attention kernels (csrc/small_mha.cu).

CPU: the oracle restatement (oracle/kine_oracle.py) against golden vectors produced by the REAL reference classes
(tests/golden/make_golden_kine.py), the classes' signatures against the reference's, name-for-name state dicts.
GPU: the attention kernels against plain fp64 / fp32 math, the modules against the same golden vectors (outputs +
gradients), autocast, CUDA-graph capture.  Tolerances: fp32 2e-5 of the tensor's max on outputs, 1e-4 on gradients
(those of tests/test_layers_gpu.py for the deformable layers); 16-bit 1.2e-2.
"""
import os
import zlib

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


def _z():
    with np.load(os.path.join(HERE, "golden", "kine_f32.npz")) as z:
        return {k: z[k] for k in z.files}


def _rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def _by_name(model):
    """tests/golden/make_golden_transformer.py::randomise_by_name."""
    with torch.no_grad():
        for name, p in model.named_parameters():
            gn = torch.Generator().manual_seed(zlib.crc32(name.encode()))
            if "norm" in name and name.endswith("weight"):
                p.copy_(1 + 0.1 * torch.randn(p.shape, generator=gn))
            elif p.dim() == 1:
                p.copy_(0.1 * torch.randn(p.shape, generator=gn))
            else:
                p.copy_(torch.randn(p.shape, generator=gn) * (0.5 / p.shape[-1] ** 0.5))
    return model


# ---------------------------------------------------------------- CPU ----------------------------------------------------------------

def test_kine_oracle_matches_reference_golden():
    from oracle import kine_oracle as ko
    z = _z()
    t = lambda k: torch.from_numpy(z[k])
    p = {k[4:]: t(k).requires_grad_(True) for k in z if k.startswith("a:p:")}
    ins = [t("a:" + k) for k in ("src_b", "src_m", "mask", "qe_b", "qe_m", "tgt_b", "tgt_m", "pos_b", "pos_m")]
    for i in (0, 1, 3, 4):
        ins[i].requires_grad_(True)
    hs_d, hs_m, raw_d, mem_d = ko.dual_kinematic_transformer(p, *ins, nhead=8, enc_layers=2, dec_layers=2)
    for got, k in ((hs_d, "hs_det"), (hs_m, "hs_meta"), (raw_d, "raw_det"), (mem_d, "mem_det")):
        assert _rel(got, t("a:" + k)) < 1e-5, k
    loss = (hs_d * t("a:w1")).sum() + (hs_m * t("a:w2")).sum() + (mem_d * t("a:w3")).sum() + (raw_d * t("a:w1")).sum() * 0.5
    names = [k[4:] for k in z if k.startswith("a:g:")]
    grads = torch.autograd.grad(loss, [ins[0], ins[1], ins[3], ins[4]] + [p[k] for k in names])
    for k, g in zip(["a:g_src_b", "a:g_src_m", "a:g_qe_b", "a:g_qe_m"] + ["a:g:" + k for k in names], grads):
        assert _rel(g, t(k)) < 2e-4, k

    # case c (encoder-only) and d (pre-norm): parameters are a function of their names
    from kinet_b200.modules import kinematic_transformer as kt
    enc = _by_name(kt.DualKinematicEncoder(d_model=72, nhead=8, num_encoder_layers=2, dim_feedforward=48, dropout=0.0))
    out = ko.dual_kinematic_encoder(dict(enc.named_parameters()), t("c:src_b"), t("c:src_m"), t("c:mask"), t("c:pos_b"),
                                    t("c:pos_m"), 8, 2)
    for got, k in zip(out, ("hs_det", "hs_meta", "mem_meta", "mem_det")):
        assert _rel(got, t("c:" + k)) < 1e-5, k
    pre = _by_name(kt.KinematicTransformer(d_model=72, nhead=8, num_encoder_layers=1, num_decoder_layers=2,
                                           dim_feedforward=48, dropout=0.0, normalize_before=True))
    out = ko.kinematic_transformer(dict(pre.named_parameters()), "", t("d:src"), t("d:mask"), t("d:qe"), t("d:tgt"),
                                   t("d:pos"), 8, 1, 2, pre_norm=True, return_intermediate=False)
    for got, k in zip(out, ("hs", "raw", "memory")):
        assert _rel(got, t("d:" + k)) < 1e-5, k


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="needs the reference checkout (build container)")
def test_kine_classes_mirror_the_reference():
    """Same class names, constructor / forward signatures and state_dict keys as the reference's transformer.py."""
    import importlib.util
    import inspect
    spec = importlib.util.spec_from_file_location("ref_kine_transformer",
                                                  "/root/reference/src/trackformer/models/transformer.py")
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    from kinet_b200.modules import kinematic_transformer as kt
    for name in ("KinematicTransformer", "DualKinematicTransformer", "DualKinematicEncoder", "TransformerEncoder",
                 "TransformerDecoder", "TransformerEncoderLayer", "TransformerDecoderLayer", "IntertwinedBranch"):
        a, b = getattr(kt, name), getattr(ref, name)
        assert list(inspect.signature(a.__init__).parameters) == list(inspect.signature(b.__init__).parameters), name
        assert list(inspect.signature(a.forward).parameters) == list(inspect.signature(b.forward).parameters), name
    kw = dict(d_model=72, nhead=8, num_encoder_layers=2, num_decoder_layers=2, dim_feedforward=48, dropout=0.0,
              return_intermediate_dec=True, track_attention=True)
    ours, theirs = kt.DualKinematicTransformer(**kw).state_dict(), ref.DualKinematicTransformer(**kw).state_dict()
    assert list(ours) == list(theirs)
    assert all(ours[k].shape == theirs[k].shape for k in ours)
    kw = dict(d_model=72, nhead=8, num_encoder_layers=2, dim_feedforward=48)
    assert list(kt.DualKinematicEncoder(**kw).state_dict()) == list(ref.DualKinematicEncoder(**kw).state_dict())


def test_kine_cpu_tensors_raise():
    from kinet_b200.functions.small_mha import small_mha
    q = torch.zeros(1, 3, 2, 4)
    with pytest.raises(RuntimeError, match="CPU"):
        small_mha(q, q, q)


def test_small_mha_argument_validation_without_gpu():
    import ctypes
    from kinet_b200 import _lib
    lib = _lib.load()
    assert lib.kinet_b200_small_mha_forward(None, None) != 0
    a = _lib.KinetMhaArgs()
    assert lib.kinet_b200_small_mha_forward(ctypes.byref(a), None) != 0          # null tensors
    a = _lib.KinetMhaArgs(8, 8, 8, 0, 0, 0, 0, 0, 0, None, 8, 0, 0, None, _lib.MSDA_F32, 1, 8, 65, 4, 4, 1.0)
    assert lib.kinet_b200_small_mha_forward(ctypes.byref(a), None) == _lib.MSDA_E_SHAPE     # D beyond KINET_MHA_MAX_D
    a.D, a.dtype = 36, _lib.MSDA_F64
    assert lib.kinet_b200_small_mha_forward(ctypes.byref(a), None) != 0          # no fp64
    b = _lib.KinetMhaBwdArgs()
    assert lib.kinet_b200_small_mha_backward(ctypes.byref(b), None) != 0


# ---------------------------------------------------------------- GPU ----------------------------------------------------------------

def _dev():
    return torch.device("cuda:0")


def _ref_attention(q, k, v, mask, scale):
    """(N, L, H, D) inputs, plain math in the inputs' dtype."""
    s = torch.einsum("nqhd,nkhd->nhqk", q, k) * scale
    if mask is not None:
        s = s.masked_fill(mask[:, None, None, :], float("-inf"))
    return torch.einsum("nhqk,nkhd->nqhd", s.softmax(-1), v).flatten(2)


@pytest.mark.gpu
@pytest.mark.parametrize("N,H,D,Lq,Lk,masked", [(2, 8, 36, 160, 23, True), (1, 8, 36, 150, 150, False),
                                                 (3, 4, 9, 7, 11, True), (2, 2, 64, 33, 130, True),
                                                 (1, 1, 1, 1, 1, False), (2, 8, 32, 17, 64, True)])
def test_small_mha_matches_fp64_math(N, H, D, Lq, Lk, masked):
    from kinet_b200.functions.small_mha import small_mha
    dev = _dev()
    g = torch.Generator().manual_seed(N * 1000 + Lq)
    # q, k live in a packed (N, L, 2 * H * D) projection output (strided views), v is sequence-first storage
    qk = torch.randn(N, max(Lq, Lk), 2 * H * D, generator=g).to(dev)
    q = qk[:, :Lq, :H * D].unflatten(-1, (H, D)).detach().requires_grad_(True)
    k = qk[:, :Lk, H * D:].unflatten(-1, (H, D)).detach().requires_grad_(True)
    assert not q.is_contiguous() or Lq == 1 or H * D == 1
    v = torch.randn(Lk, N, H, D, generator=g).to(dev).transpose(0, 1).detach().requires_grad_(True)
    mask = None
    if masked:
        mask = torch.rand(N, Lk, generator=g) < 0.3
        mask[:, 0] = False                              # keep one live key per row
        mask = mask.to(dev)
    go = torch.randn(N, Lq, H * D, generator=g).to(dev)
    out = small_mha(q, k, v, mask)
    gq, gk, gv = torch.autograd.grad(out, [q, k, v], go)
    qd, kd, vd = (t.detach().double().requires_grad_(True) for t in (q, k, v))
    ref = _ref_attention(qd, kd, vd, mask, D ** -0.5)
    rq, rk, rv = torch.autograd.grad(ref, [qd, kd, vd], go.double())
    for got, want, name in ((out, ref, "out"), (gq, rq, "dq"), (gk, rk, "dk"), (gv, rv, "dv")):
        assert _rel(got, want) < 2e-5, name
    # bit-reproducible (one writer per element, fixed summation order)
    out2 = small_mha(q, k, v, mask)
    assert torch.equal(out, out2)
    assert all(torch.equal(a, b) for a, b in zip((gq, gk, gv), torch.autograd.grad(out2, [q, k, v], go)))


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_small_mha_16bit(dtype):
    from kinet_b200.functions.small_mha import small_mha
    dev = _dev()
    g = torch.Generator().manual_seed(5)
    N, H, D, Lq, Lk = 2, 8, 36, 50, 70
    q, k, v = (torch.randn(N, L, H, D, generator=g).to(dev, dtype).requires_grad_(True) for L in (Lq, Lk, Lk))
    mask = (torch.rand(N, Lk, generator=g) < 0.2).to(dev)
    mask[:, 3] = False
    go = torch.randn(N, Lq, H * D, generator=g).to(dev, dtype)
    out = small_mha(q, k, v, mask)
    grads = torch.autograd.grad(out, [q, k, v], go)
    qd, kd, vd = (t.detach().double().requires_grad_(True) for t in (q, k, v))
    ref = _ref_attention(qd, kd, vd, mask, D ** -0.5)
    rgrads = torch.autograd.grad(ref, [qd, kd, vd], go.double())
    assert out.dtype == dtype
    for got, want in zip((out,) + grads, (ref,) + rgrads):
        assert _rel(got, want) < 1.2e-2


@pytest.mark.gpu
def test_small_mha_all_keys_masked_rows_are_zero_like_torch():
    """torch >= 2.5 returns 0 for a query whose keys are all masked (safe softmax); torch 1.5 returned NaN."""
    from kinet_b200.functions.small_mha import small_mha
    dev = _dev()
    q = torch.randn(2, 5, 2, 8, device=dev, requires_grad=True)
    mask = torch.zeros(2, 5, dtype=torch.bool, device=dev)
    mask[1] = True
    out = small_mha(q, q, q, mask)
    assert torch.isfinite(out).all() and out[0].abs().max() > 0 and not out[1].any()
    g, = torch.autograd.grad(out.sum(), q)
    assert torch.isfinite(g).all() and not g[1].any()
    mha = torch.nn.MultiheadAttention(16, 2).to(dev).eval()
    x = torch.randn(5, 2, 16, device=dev)
    with torch.no_grad():
        want = mha(x, x, x, key_padding_mask=mask, need_weights=False)[0]
    assert torch.allclose(want[:, 1], mha.out_proj.bias.expand(5, 16), atol=1e-6)     # zero attention output


def _load_a(dev):
    from kinet_b200.modules import kinematic_transformer as kt
    z = _z()
    model = kt.DualKinematicTransformer(d_model=72, nhead=8, num_encoder_layers=2, num_decoder_layers=2,
                                        dim_feedforward=96, dropout=0.0, return_intermediate_dec=True)
    sd = {k[4:]: torch.from_numpy(v) for k, v in z.items() if k.startswith("a:p:")}
    missing, unexpected = model.load_state_dict(sd, strict=True)
    assert not missing and not unexpected
    t = lambda k: torch.from_numpy(z["a:" + k]).to(dev)
    ins = [t(k) for k in ("src_b", "src_m", "mask", "qe_b", "qe_m", "tgt_b", "tgt_m", "pos_b", "pos_m")]
    return z, model.to(dev).eval(), ins, t


@pytest.mark.gpu
def test_dual_kinematic_transformer_matches_reference_golden_fp32():
    dev = _dev()
    z, model, ins, t = _load_a(dev)
    for i in (0, 1, 3, 4):
        ins[i].requires_grad_(True)
    hs_d, hs_m, raw_d, mem_d = model(*ins)
    for got, k in ((hs_d, "hs_det"), (hs_m, "hs_meta"), (raw_d, "raw_det"), (mem_d, "mem_det")):
        assert got.shape == t(k).shape, k
        assert _rel(got, t(k)) < 2e-5, k
    loss = (hs_d * t("w1")).sum() + (hs_m * t("w2")).sum() + (mem_d * t("w3")).sum() + (raw_d * t("w1")).sum() * 0.5
    params = dict(model.named_parameters())
    names = [k[4:] for k in z if k.startswith("a:g:")]
    assert len(names) >= 80
    grads = torch.autograd.grad(loss, [ins[0], ins[1], ins[3], ins[4]] + [params[k] for k in names])
    for k, g in zip(["g_src_b", "g_src_m", "g_qe_b", "g_qe_m"] + ["g:" + k for k in names], grads):
        assert _rel(g, t(k)) < 1e-4, k


@pytest.mark.gpu
def test_kine_configuration_width_encoder_only_and_pre_norm_match_golden():
    """case b: d_model 288 = 8 heads of 36 channels, 160 queries (cfgs/train_kinet.yaml); c: DualKinematicEncoder;
    d: pre-norm KinematicTransformer through the reference's sequence-first interface."""
    from kinet_b200.modules import kinematic_transformer as kt
    dev = _dev()
    z = _z()
    t = lambda k: torch.from_numpy(z[k]).to(dev)
    model = _by_name(kt.DualKinematicTransformer(d_model=288, nhead=8, num_encoder_layers=1, num_decoder_layers=1,
                                                 dim_feedforward=64, dropout=0.0, return_intermediate_dec=True)).to(dev).eval()
    with torch.no_grad():
        out = model(*[t("b:" + k) for k in ("src_b", "src_m", "mask", "qe_b", "qe_m", "tgt_b", "tgt_m", "pos_b", "pos_m")])
    for got, k in zip(out, ("hs_det", "hs_meta", "raw_det", "mem_det")):
        assert got.shape == t("b:" + k).shape and _rel(got, t("b:" + k)) < 2e-5, k
    enc = _by_name(kt.DualKinematicEncoder(d_model=72, nhead=8, num_encoder_layers=2, dim_feedforward=48, dropout=0.0)).to(dev).eval()
    with torch.no_grad():
        out = enc(t("c:src_b"), t("c:src_m"), t("c:mask"), t("c:pos_b"), t("c:pos_m"))
    for got, k in zip(out, ("hs_det", "hs_meta", "mem_meta", "mem_det")):
        assert got.shape == t("c:" + k).shape and _rel(got, t("c:" + k)) < 2e-5, k
    pre = _by_name(kt.KinematicTransformer(d_model=72, nhead=8, num_encoder_layers=1, num_decoder_layers=2,
                                           dim_feedforward=48, dropout=0.0, normalize_before=True)).to(dev).eval()
    with torch.no_grad():
        out = pre(t("d:src"), t("d:mask"), t("d:qe"), t("d:tgt"), t("d:pos"))
    for got, k in zip(out, ("hs", "raw", "memory")):
        assert got.shape == t("d:" + k).shape and _rel(got, t("d:" + k)) < 2e-5, k


@pytest.mark.gpu
def test_kine_layers_sequence_first_interface_matches_torch_modules():
    """A single encoder / decoder layer called the way the reference calls it (sequence-first tensors), against the
    same arithmetic written with nn.MultiheadAttention."""
    from kinet_b200.modules import kinematic_transformer as kt
    from oracle import kine_oracle as ko
    dev = _dev()
    g = torch.Generator().manual_seed(9)
    d, N, T, Lq = 72, 2, 10, 6
    enc = _by_name(kt.TransformerEncoderLayer(d, 8, 48, 0.0)).to(dev).eval()
    dec = _by_name(kt.TransformerDecoderLayer(d, 8, 48, 0.0)).to(dev).eval()
    src, pos = torch.randn(T, N, d, generator=g).to(dev), torch.randn(T, N, d, generator=g).to(dev)
    tgt, qpos = torch.randn(Lq, N, d, generator=g).to(dev), torch.randn(Lq, N, d, generator=g).to(dev)
    mask = torch.zeros(N, T, dtype=torch.bool, device=dev)
    mask[0, -2:] = True
    with torch.no_grad():
        mem = enc(src, src_key_padding_mask=mask, pos=pos)
        out = dec(tgt, mem, memory_key_padding_mask=mask, pos=pos, query_pos=qpos)
        want_mem = ko.encoder_layer(dict(enc.named_parameters()), "", src, mask, pos, 8, False)
        want_out = ko.decoder_layer(dict(dec.named_parameters()), "", tgt, want_mem, mask, pos, qpos, 8, False)
    assert mem.shape == src.shape and out.shape == tgt.shape
    assert _rel(mem, want_mem) < 2e-5 and _rel(out, want_out) < 2e-5
    with pytest.raises(RuntimeError, match="attn_mask"):
        enc(src, src_mask=torch.zeros(T, T, device=dev))


@pytest.mark.gpu
def test_kine_autocast_and_cuda_graph():
    """bf16 autocast stays within the 16-bit bar of the fp32 result, and the whole forward records into a CUDA graph
    (no allocation-free requirement here, but no host sync / host read anywhere on the path)."""
    dev = _dev()
    z, model, ins, t = _load_a(dev)
    with torch.no_grad():
        want = model(*ins)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            got = model(*ins)
    for a, b in zip(got, want):
        assert a.dtype == torch.float32          # the residual stream keeps its dtype
        assert _rel(a, b) < 3e-2
    static = [x.clone() for x in ins]
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s), torch.no_grad():
        for _ in range(2):
            model(*static)
    torch.cuda.current_stream().wait_stream(s)
    graph = torch.cuda.CUDAGraph()
    with torch.no_grad(), torch.cuda.graph(graph):
        out = model(*static)
    static[0].copy_(ins[0] * 0.5)
    graph.replay()
    with torch.no_grad():
        want2 = model(*([ins[0] * 0.5] + ins[1:]))
    for a, b in zip(out, want2):
        assert torch.equal(a, b)


@pytest.mark.gpu
def test_kine_track_attention_and_training_mode():
    """track_attention=True (the last 100 queries are object queries, :290-305) against the same composition written
    with the oracle; dropout > 0 in training mode runs (library attention with dropout_p) and backpropagates."""
    from kinet_b200.modules import kinematic_transformer as kt
    from oracle import kine_oracle as ko
    dev = _dev()
    g = torch.Generator().manual_seed(4)
    d, N, T, Lq = 72, 2, 9, 104
    model = _by_name(kt.KinematicTransformer(d_model=d, nhead=8, num_encoder_layers=1, num_decoder_layers=2,
                                             dim_feedforward=48, dropout=0.0, return_intermediate_dec=True,
                                             track_attention=True)).to(dev).eval()
    src, pos = torch.randn(N, T, d, generator=g).to(dev), torch.randn(N, T, d, generator=g).to(dev)
    qe, tgt = torch.randn(Lq, N, d, generator=g).to(dev), torch.randn(Lq, N, d, generator=g).to(dev)
    qe0 = qe.clone()
    with torch.no_grad():
        hs, raw, memory = model(src, None, qe, tgt, pos)
    assert torch.equal(qe, qe0)                   # (the reference zeroes the caller's query_pos in place; we do not)
    p = dict(model.named_parameters())
    with torch.no_grad():
        mem = ko.encoder(p, "encoder.", src.permute(1, 0, 2), None, pos.permute(1, 0, 2), 8, 1, False)
        tq = qe[:-100].clone()
        qp = qe.clone()
        qp[:-100] = 0
        out, inter = tgt, []
        for i in range(2):
            tr = ko.encoder_layer(p, "decoder.layers_track_attention.%d." % i, out[:-100], None, tq, 8, False)
            out = torch.cat([tr, out[-100:]])
            out = ko.decoder_layer(p, "decoder.layers.%d." % i, out, mem, None, pos.permute(1, 0, 2), qp, 8, False)
            inter.append(out)
        want_raw = torch.stack(inter)
    assert _rel(memory, mem) < 2e-5 and _rel(raw, want_raw) < 2e-5
    train = kt.DualKinematicTransformer(d_model=d, nhead=8, num_encoder_layers=1, num_decoder_layers=1,
                                        dim_feedforward=48, dropout=0.1, return_intermediate_dec=True).to(dev).train()
    src_r = src.clone().requires_grad_(True)
    outs = train(src_r, src, None, qe[:7], qe[:7], None, None, pos, pos)
    sum(o.float().sum() for o in outs).backward()
    assert torch.isfinite(src_r.grad).all() and src_r.grad.abs().max() > 0


def _detector(dev, Lq=6, d=72):
    from kinet_b200.modules import kinematic_transformer as kt
    from kinet_b200.modules.kinematic_detector import KinematicDetectorTransformer, TokenFeatures
    holder = {}
    back = [lambda x: ([TokenFeatures(x[0], holder["mask"])], [x[1]])] * 2
    transformer = kt.DualKinematicTransformer(d_model=d, nhead=8, num_encoder_layers=1, num_decoder_layers=2,
                                              dim_feedforward=48, dropout=0.0, return_intermediate_dec=True)
    model = KinematicDetectorTransformer(back, transformer, num_classes=1, num_queries=Lq, aux_loss=True,
                                         dim_tracklets_det=16, dim_tracklets_metadata=8)
    return _by_name(model).to(dev).eval(), holder


class _Samples:
    def __init__(self, det, meta):
        self.detections, self.metadata = det, meta


@pytest.mark.gpu
def test_kinematic_detector_matches_reference_golden_and_takes_track_queries():
    """case e: the reference's KinematicDetectorTransformer (detr.py:288-417) without track queries; with track
    queries (``targets``) the reference's own glue fails (see the module), so that path is checked against the
    oracle composition of the intended arithmetic."""
    from oracle import kine_oracle as ko
    dev = _dev()
    z = _z()
    t = lambda k: torch.from_numpy(z["e:" + k]).to(dev)
    model, holder = _detector(dev)
    holder["mask"] = t("mask")
    smp = _Samples((t("src_b"), t("pos_b")), (t("src_m"), t("pos_m")))
    with torch.no_grad():
        out, targets, feats, src_out, hs_det = model(smp, None)
    assert targets is None and torch.equal(src_out, t("src_out"))
    for got, k in ((out["pred_logits"], "pred_logits"), (out["pred_boxes"], "pred_boxes"),
                   (out["aux_outputs"][0]["pred_logits"], "aux_logits"), (out["aux_outputs"][0]["pred_boxes"], "aux_boxes"),
                   (hs_det, "hs_det")):
        assert got.shape == t(k).shape and _rel(got, t(k)) < 2e-5, k
    # 3 track queries per sample
    g = torch.Generator().manual_seed(12)
    tg = [{"track_query_hs_embeds_det": torch.randn(3, 16, generator=g).to(dev),
           "track_query_hs_embeds_meta": torch.randn(3, 8, generator=g).to(dev)} for _ in range(2)]
    with torch.no_grad():
        out, _, _, _, hs_det = model(smp, tg)
        p = dict(model.named_parameters())
        mlp = lambda pre, x: ko._lin(p, pre + ".layers.2", torch.relu(ko._lin(p, pre + ".layers.1", torch.relu(ko._lin(p, pre + ".layers.0", x)))))
        N, Lq, d = 2, 6, 72
        qd = torch.cat([torch.zeros(3, N, d, device=dev), p["query_embed_det.weight"][:, None].repeat(1, N, 1)])
        qm = torch.cat([torch.zeros(3, N, d, device=dev), p["query_embed_metadata.weight"][:, None].repeat(1, N, 1)])
        td, tm = torch.zeros_like(qd), torch.zeros_like(qm)
        td[:3] = mlp("input_proj_tracklets_det", torch.stack([x["track_query_hs_embeds_det"] for x in tg]).transpose(0, 1))
        tm[:3] = mlp("input_proj_tracklets_metadata", torch.stack([x["track_query_hs_embeds_meta"] for x in tg]).transpose(0, 1))
        pt = {k[len("transformer."):]: v for k, v in p.items() if k.startswith("transformer.")}
        hd, hm, _, _ = ko.dual_kinematic_transformer(pt, t("src_b"), t("src_m"), t("mask"), qd, qm, td, tm, t("pos_b"),
                                                      t("pos_m"), 8, 1, 2)
        want_logits = ko._lin(p, "class_embed", hm)[-1]
        want_boxes = mlp("bbox_embed", hd).sigmoid()[-1]
    assert out["pred_logits"].shape == (2, 9, 2) and out["pred_boxes"].shape == (2, 9, 4)
    assert _rel(out["pred_logits"], want_logits) < 2e-5 and _rel(out["pred_boxes"], want_boxes) < 2e-5
    assert _rel(hs_det, hd) < 2e-5


@pytest.mark.gpu
def test_kine_frame_graphs_match_direct_calls_with_varying_sizes():
    """KineFrameGraphs: token counts padded to a bucket and masked, track queries at their own count (track_bucket = 1):
    every frame equals the direct call; the graphs are recorded once per (padded tokens, track queries) pair."""
    import torch.nn.functional as F
    from kinet_b200.frame_graph import KineFrameGraphs
    from kinet_b200.modules import kinematic_transformer as kt
    dev = _dev()
    d, nq = 72, 6
    model = _by_name(kt.DualKinematicTransformer(d_model=d, nhead=8, num_encoder_layers=1, num_decoder_layers=1,
                                                 dim_feedforward=48, dropout=0.0, return_intermediate_dec=True)).to(dev).eval()
    g = torch.Generator().manual_seed(8)
    qe_d, qe_m = torch.randn(nq, 1, d, generator=g).to(dev), torch.randn(nq, 1, d, generator=g).to(dev)

    def heads(transformer, bufs, Kp):
        qd, qm = F.pad(qe_d, (0, 0, 0, 0, Kp, 0)), F.pad(qe_m, (0, 0, 0, 0, Kp, 0))
        td = F.pad(bufs["tgt_det"], (0, 0, 0, 0, 0, nq)) if Kp else None
        tm = F.pad(bufs["tgt_meta"], (0, 0, 0, 0, 0, nq)) if Kp else None
        qpad = F.pad(bufs["kpad"], (0, nq)) if Kp else None
        hs_det, hs_meta, _, _ = transformer.forward_masked_queries(bufs["src_det"], bufs["src_meta"], bufs["mask"], qd, qm,
                                                                   td, tm, bufs["pos_det"], bufs["pos_meta"], qpad)
        return hs_det[-1], hs_meta[-1]
    fg = KineFrameGraphs(model, heads, token_bucket=8, track_bucket=4, autocast_dtype=None)
    sizes = [(5, 0), (11, 3), (8, 3), (13, 2), (11, 4), (6, 0), (9, 5)]
    for T, K in sizes:
        toks = [torch.randn(1, T, d, generator=g).to(dev) for _ in range(4)]
        tg = [torch.randn(K, 1, d, generator=g).to(dev) for _ in range(2)] if K else [None, None]
        (hd, hm), pad = fg(*toks, *tg)
        assert pad == -K % 4
        hd, hm = hd[:, pad:], hm[:, pad:]               # padding track-query rows sit at the front
        with torch.no_grad():
            qd, qm = F.pad(qe_d, (0, 0, 0, 0, K, 0)), F.pad(qe_m, (0, 0, 0, 0, K, 0))
            td = F.pad(tg[0], (0, 0, 0, 0, 0, nq)) if K else None
            tm = F.pad(tg[1], (0, 0, 0, 0, 0, nq)) if K else None
            want_d, want_m, _, _ = model(toks[0], toks[1], torch.zeros(1, T, dtype=torch.bool, device=dev), qd, qm, td, tm,
                                         toks[2], toks[3])
        assert hd.shape == want_d[-1].shape
        assert _rel(hd, want_d[-1]) < 2e-5 and _rel(hm, want_m[-1]) < 2e-5, (T, K)
    assert fg.recorded_buckets == [(8, 0), (8, 4), (16, 4), (16, 8)]
    assert fg.replays == len(sizes)


@pytest.mark.gpu
@pytest.mark.parametrize("patience", [-1, 3])
def test_kine_sequence_tracker_graphs_match_eager_detector_over_a_sequence(patience):
    """KineSequenceTracker (the loop of TrackerKinematic.step: trails -> track queries -> detector -> thresholds / NMS ->
    births and terminations): replaying KineFrameGraphs gives the same track ids every frame and boxes within 1e-3 px of the
    eager loop through KinematicDetectorTransformer.forward(samples, targets); tracks are born, carried and terminated."""
    from kinet_b200.modules import kinematic_transformer as kt
    from kinet_b200.modules.kinematic_detector import KinematicDetectorTransformer, TokenFeatures
    from kinet_b200.tracking import KineSequenceTracker
    dev = _dev()
    d, nq, n_frames = 72, 12, 5
    back = [lambda x: ([TokenFeatures(x[0], torch.zeros(x[0].shape[:2], dtype=torch.bool, device=x[0].device))], [x[1]])] * 2
    transformer = kt.DualKinematicTransformer(d_model=d, nhead=8, num_encoder_layers=1, num_decoder_layers=1,
                                              dim_feedforward=48, dropout=0.0, return_intermediate_dec=True)
    det = _by_name(KinematicDetectorTransformer(back, transformer, num_classes=1, num_queries=nq, aux_loss=False,
                                                dim_tracklets_det=4 * n_frames, dim_tracklets_metadata=n_frames)).to(dev).eval()
    with torch.no_grad():
        det.class_embed.weight.mul_(6.0)                      # spread the scores around the thresholds
        det.bbox_embed.layers[-1].weight.mul_(8.0)            # ... and the boxes over the image (NMS would merge them)
    g = torch.Generator().manual_seed(21)
    scene = [torch.randn(1, 20, d, generator=g) for _ in range(4)]
    frames = []
    for i in range(24):
        T = 9 + (7 * i) % 11
        frames.append(tuple((s[:, :T] + 0.25 * torch.randn(1, T, d, generator=g)).to(dev) for s in scene))
    # patience 3: terminated tracks stay inactive track queries for three frames (re-identification, cfgs/track_kinet.yaml:45)
    kw = dict(img_size=(1080, 1920), n_frames=n_frames, detection_nms_thresh=0.9, track_nms_thresh=0.9,
              inactive_patience=patience, reid_score_thresh=0.0)
    a = KineSequenceTracker(det, use_graphs=True, token_bucket=8, track_bucket=4, **kw)
    b = KineSequenceTracker(det, use_graphs=False, **kw)
    # random weights: put the thresholds where about a third of the object queries fire
    with torch.no_grad():
        sc = b._detect_eager(frames[0], None, None)[:, 0]
    thr = float(torch.quantile(sc, 0.67))
    for t in (a, b):
        t.det_thresh, t.trk_thresh = thr, thr - 0.02 * (float(sc.max()) - float(sc.min()))
        t.reid_thresh = t.trk_thresh                       # an inactive track comes back when it scores like a live one
    counts, inactive_seen = [], False
    trk_thresh = a.trk_thresh
    for i, f in enumerate(frames):
        if patience >= 0:
            # random weights give near-binary scores: drive the inactive path by hand -- every 6th frame all live tracks
            # are terminated (they stay inactive track queries), two frames later every inactive query is taken back
            for t in (a, b):
                t.trk_thresh = float("inf") if i % 6 == 3 else trk_thresh
                t.reid_thresh = float("-inf") if i % 6 == 5 else float("inf")
        ra, rb = a.step(f), b.step(f)
        assert sorted(ra) == sorted(rb)
        for tid in ra:
            assert np.abs(ra[tid]["bbox"] - rb[tid]["bbox"]).max() < 1e-3
            assert abs(ra[tid]["score"] - rb[tid]["score"]) < 1e-5
        assert a.inactive_tracks == b.inactive_tracks
        counts.append(len(ra))
        inactive_seen = inactive_seen or bool(a.inactive_tracks)
    assert inactive_seen == (patience >= 0) and a.num_reids == b.num_reids and (a.num_reids > 0) == (patience >= 0)
    assert a.next_id == b.next_id and a.next_id > max(counts) > 0          # tracks were started and some ended
    assert any(len(v) >= 3 for v in a.results.values())                   # some track lived for several frames
    assert len(a._graphs.recorded_buckets) >= 3
