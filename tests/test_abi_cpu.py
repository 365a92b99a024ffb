"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/msda_b200.h declares, and refuses to compute without a CUDA device (no CPU fallback)."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from kinet_b200 import build
    build.build()
    from kinet_b200 import _lib
    return _lib.load()


def test_header_symbols_are_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "msda_b200.h")).read()
    declared = set(re.findall(r"\b(msda_b200_\w+)\s*\(", hdr))
    assert {"msda_b200_forward", "msda_b200_backward", "msda_b200_backward_workspace_bytes",
            "msda_b200_error_string", "msda_b200_abi_version",
            "msda_b200_last_launch_count", "msda_b200_trace_next"} <= declared
    for name in declared:
        assert hasattr(lib, name), name
    from kinet_b200 import _lib
    assert set(_lib.EXPORTS) == declared
    hdr2 = open(os.path.join(ROOT, "include", "kinet_layers_b200.h")).read()
    declared2 = set(re.findall(r"\b(kinet_b200_\w+)\s*\(", hdr2))
    assert len(declared2) == 9
    for name in declared2:
        assert hasattr(lib, name), name
    assert set(_lib.LAYER_EXPORTS) == declared2


def test_abi_version_and_error_strings(lib):
    assert lib.msda_b200_abi_version() == 3
    assert lib.msda_b200_error_string(0) == b"success"
    for code in (-1, -2, -3, -4, -5):
        assert lib.msda_b200_error_string(code).startswith(b"msda_b200:")


def test_workspace_size(lib):
    from kinet_b200._lib import MsdaDims, MSDA_BF16, MSDA_F32
    d = MsdaDims(2, 100, 8, 32, 4, 7, 4)
    assert lib.msda_b200_backward_workspace_bytes(MSDA_F32, ctypes.byref(d)) == 0
    # fp32 accumulator (rounded up to 256 bytes) + the dense-window kernel's work counter (256 bytes) and its
    # done-item flags: one byte per (tile, batch, head), at most Lq / 32 + 64 tiles
    acc = 2 * 100 * 8 * 32 * 4
    flags = (7 // 32 + 64) * 2 * 8
    assert lib.msda_b200_backward_workspace_bytes(MSDA_BF16, ctypes.byref(d)) == (acc + 255) // 256 * 256 + 256 + (flags + 255) // 256 * 256
    # sparse query sets (4 * Lq * L * P <= S: the decoder form) accumulate in the 16-bit output: no workspace
    sparse = MsdaDims(2, 22223, 8, 32, 4, 300, 4)
    assert lib.msda_b200_backward_workspace_bytes(MSDA_BF16, ctypes.byref(sparse)) == 0


def test_argument_validation_without_gpu(lib):
    from kinet_b200._lib import MsdaDims
    d = MsdaDims(2, 100, 8, 32, 4, 7, 4)
    null = ctypes.c_void_p(0)
    assert lib.msda_b200_forward(null, 0, null, null, null, null, 0, null, ctypes.byref(d), null) == -3
    bad = MsdaDims(2, 100, 8, 32, 99, 7, 4)
    one = ctypes.c_void_p(16)
    assert lib.msda_b200_forward(one, 0, one, null, one, one, 0, one, ctypes.byref(bad), null) == -2
    assert lib.msda_b200_forward(one, 7, one, null, one, one, 0, one, ctypes.byref(d), null) == -1


def test_layer_abi_argument_validation_without_gpu(lib):
    """include/kinet_layers_b200.h entry points check their arguments before touching CUDA."""
    from kinet_b200._lib import KinetLnBwdArgs, KinetLnFwdArgs, MSDA_BF16, MSDA_F32
    p16, p8, p2 = 4096, 4104, 4098          # fake device addresses: 16-, 8- and 2-byte aligned
    assert lib.kinet_b200_add_layernorm_forward(None, None) == -3

    def fwd(**kw):
        base = dict(x=p16, x_dtype=MSDA_F32, r=p8, r_dtype=MSDA_BF16, pos=None, pos_dtype=0, pos_rows=1,
                    gamma=p16, beta=p16, y=p16, y_dtype=MSDA_F32, y_lp=None, ypos_lp=None, lp_dtype=MSDA_BF16,
                    mean_rstd=None, rows=0, C=256, eps=1e-5)
        base.update(kw)
        return lib.kinet_b200_add_layernorm_forward(ctypes.byref(KinetLnFwdArgs(**base)), None)
    assert fwd() == 0                                   # rows == 0: nothing to do, no launch
    assert fwd(x=None) == -3 and fwd(gamma=None) == -3
    assert fwd(C=6) == -2 and fwd(C=1028) == -2 and fwd(C=0) == -2 and fwd(rows=-1) == -2
    assert fwd(x_dtype=2) == -1 and fwd(r_dtype=7) == -1 and fwd(y_lp=p8, lp_dtype=2) == -1 and fwd(r_dtype=3) == 0
    assert fwd(x=p8) == -4 and fwd(r=p2) == -4 and fwd(y_lp=p2) == -4
    assert fwd(ypos_lp=p8) == -3                        # y + pos asked for without pos
    assert fwd(ypos_lp=p8, pos=p16, pos_rows=0) == -3

    def bwd(**kw):
        base = dict(dy=p16, dy_dtype=MSDA_F32, dy_lp=None, dypos_lp=None, lp_dtype=MSDA_BF16, x=p16, x_dtype=MSDA_F32,
                    r=p8, r_dtype=MSDA_BF16, gamma=p16, mean_rstd=p16, dx=p16, dr=p8, dgamma_dbeta=p16, partial=p16,
                    rows=4, C=256)
        base.update(kw)
        return lib.kinet_b200_add_layernorm_backward(ctypes.byref(KinetLnBwdArgs(**base)), None)
    assert bwd(dy=None) == -3 and bwd(partial=None) == -3 and bwd(r=None) == -3      # dr without r
    assert bwd(C=2) == -2 and bwd(dy_dtype=2) == -1 and bwd(dx=p8) == -4
    assert lib.kinet_b200_colsum_partial_floats(256) >= 148 * 256 and lib.kinet_b200_colsum_partial_floats(0) == 0
    assert lib.kinet_b200_colsum(p16, MSDA_BF16, 10, 256, None, MSDA_F32, p16, None) == -3
    assert lib.kinet_b200_colsum(p16, MSDA_BF16, 10, 258, p16, MSDA_F32, p16, None) == -2
    assert lib.kinet_b200_colsum(p16, 2, 10, 256, p16, MSDA_F32, p16, None) == -1
    assert lib.kinet_b200_colsum(p2, MSDA_BF16, 10, 256, p16, MSDA_F32, p16, None) == -4
    assert lib.kinet_b200_relu_backward_colsum(p16, None, MSDA_BF16, 10, 256, p16, p16, MSDA_F32, p16, None) == -3


def test_layer_modules_mirror_the_reference_signatures():
    """Constructor / forward signatures and parameter names of the reference's layer classes
    (src/trackformer/models/deformable_transformer.py:260-433), checked without a GPU."""
    import inspect
    from kinet_b200.modules import (DeformableTransformerDecoder, DeformableTransformerDecoderLayer,
                                    DeformableTransformerEncoder, DeformableTransformerEncoderLayer)
    args = lambda f: list(inspect.signature(f).parameters)
    assert args(DeformableTransformerEncoderLayer.__init__)[1:] == ["d_model", "d_ffn", "dropout", "activation",
                                                                    "n_levels", "n_heads", "n_points"]
    assert args(DeformableTransformerEncoderLayer.forward)[1:] == ["src", "pos", "reference_points", "spatial_shapes",
                                                                   "padding_mask"]
    assert args(DeformableTransformerEncoder.forward)[1:6] == ["src", "spatial_shapes", "valid_ratios", "pos",
                                                               "padding_mask"]
    assert args(DeformableTransformerDecoderLayer.forward)[1:] == ["tgt", "query_pos", "reference_points", "src",
                                                                   "src_spatial_shapes", "src_padding_mask",
                                                                   "query_attn_mask"]
    assert args(DeformableTransformerDecoder.forward)[1:] == ["tgt", "reference_points", "src", "src_spatial_shapes",
                                                              "src_valid_ratios", "query_pos", "src_padding_mask",
                                                              "query_attn_mask"]
    enc = DeformableTransformerEncoder(DeformableTransformerEncoderLayer(64, 32, 0.1, "relu", 2, 8, 2), 2)
    dec = DeformableTransformerDecoder(DeformableTransformerDecoderLayer(64, 32, 0.1, "relu", 2, 8, 2), 2, True)
    names = set(k for k, _ in enc.named_parameters())
    assert {"layers.1.self_attn.sampling_offsets.weight", "layers.0.norm2.bias", "layers.0.linear1.weight"} <= names
    names = set(k for k, _ in dec.named_parameters())
    assert {"layers.0.cross_attn.value_proj.bias", "layers.1.self_attn.in_proj_weight",
            "layers.0.self_attn.out_proj.bias", "layers.0.norm3.weight"} <= names
    assert dec.bbox_embed is None and dec.class_embed is None and dec.return_intermediate
    # reference points of the encoder: pixel centres / (valid ratio * size), for every level (CPU tensors are fine here)
    shapes = torch.tensor([[2, 3], [1, 2]])
    ref = DeformableTransformerEncoder.get_reference_points(shapes, torch.ones(1, 2, 2), torch.device("cpu"))
    assert ref.shape == (1, 8, 2, 2)
    assert torch.allclose(ref[0, :3, 0, 0], torch.tensor([0.5, 1.5, 2.5]) / 3) and torch.allclose(ref[0, 6:, 1, 0], torch.tensor([0.25, 0.75]))
    with pytest.raises(RuntimeError, match="Not implemented on the CPU"):
        enc(torch.zeros(1, 8, 64), shapes, torch.ones(1, 2, 2), torch.zeros(1, 8, 64))


def test_cpu_tensors_raise_like_the_reference():
    from kinet_b200.functions import MSDeformAttnFunction
    value = torch.zeros(1, 4, 1, 4)
    shapes = torch.tensor([[2, 2]])
    loc = torch.zeros(1, 1, 1, 1, 1, 2)
    attn = torch.ones(1, 1, 1, 1, 1)
    with pytest.raises(RuntimeError, match="Not implemented on the CPU"):
        MSDeformAttnFunction.apply(value, shapes, loc, attn, 64)


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "kinet_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f
                assert "msda_oracle" not in src, f


def test_extension_module_name_shim():
    """`import MultiScaleDeformableAttention as MSDA` (reference ms_deform_attn_func.py:11) resolves to
    the shim exporting the two pybind names of src/vision.cpp:5-6, and rejects CPU tensors."""
    import MultiScaleDeformableAttention as MSDA
    assert callable(MSDA.ms_deform_attn_forward) and callable(MSDA.ms_deform_attn_backward)
    v = torch.zeros(1, 4, 1, 4)
    shapes = torch.tensor([[2, 2]])
    loc = torch.zeros(1, 1, 1, 1, 1, 2)
    attn = torch.zeros(1, 1, 1, 1, 1)
    with pytest.raises(RuntimeError, match="Not implemented on the CPU"):
        MSDA.ms_deform_attn_forward(v, shapes, loc, attn, 64)
    with pytest.raises(TypeError):
        MSDA.ms_deform_attn_forward(v, shapes, loc)


def test_graft_entry_build_returns():
    """build() must return to its caller (an earlier version ended in sys.exit via runpy)."""
    import __graft_entry__ as g
    g.build()
    from kinet_b200 import _lib
    assert _lib.load() is not None


def test_per_tensor_caches_survive_alternating_level_lists(monkeypatch):
    """The multi-frame configuration alternates between two spatial_shapes tensors (encoder: L levels, decoder: 2L);
    the per-tensor caches must hit for both, re-evaluate after an in-place change, and never serve a new tensor
    that happens to reuse an id."""
    from kinet_b200._cache import PerTensorCache
    from kinet_b200.modules import ms_deform_attn as mod
    from kinet_b200.modules.deformable_transformer import _level_grids, _shapes_tensor
    enc, dec = torch.tensor([[2, 3], [1, 2]]), torch.tensor([[2, 3], [1, 2], [2, 3], [1, 2]])
    mod._shapes_validated = PerTensorCache()
    evaluations = []
    real_sum = torch.Tensor.sum
    monkeypatch.setattr(torch.Tensor, "sum", lambda self, *a, **k: (evaluations.append(1), real_sum(self, *a, **k))[1])
    for _ in range(4):
        mod._check_spatial_shapes(enc, 8)
        mod._check_spatial_shapes(dec, 16)
    assert len(evaluations) == 2                       # once per tensor object, not once per call
    with pytest.raises(AssertionError):
        mod._check_spatial_shapes(enc, 9)              # another Len_in is another question
    enc[0, 0] = 5                                      # in-place change: version counter moves, cache misses
    with pytest.raises(AssertionError):
        mod._check_spatial_shapes(enc, 8)
    monkeypatch.undo()
    g1 = _level_grids(dec, torch.device("cpu"))
    assert _level_grids(dec, torch.device("cpu")) is g1 and len(g1) == 4 and g1[0][0].numel() == 6
    assert _shapes_tensor([(2, 3), (1, 2)], "cpu") is _shapes_tensor([(2, 3), (1, 2)], "cpu")
    c = PerTensorCache(maxsize=2)
    ts = [torch.zeros(1) for _ in range(3)]
    for i, t in enumerate(ts):
        c.put(t, i)
    assert len(c) == 2 and c.get(ts[0]) is None and c.get(ts[2]) == 2
    t = torch.zeros(1)
    c.put(t, "old")
    key = id(t)
    del t
    t2 = torch.ones(1)                                 # may or may not reuse the id; a dead weakref must never hit
    assert c.get(t2) is None or id(t2) != key


def test_per_tensor_cache_accepts_inference_tensors():
    """Tensors created under torch.inference_mode() have no version counter (reading it raises); the reference
    module works there, so must ours (the shapes tensor is built inside forward, like the reference's)."""
    from kinet_b200._cache import PerTensorCache
    from kinet_b200.modules import ms_deform_attn as mod
    from kinet_b200.modules.deformable_transformer import _level_grids
    with torch.inference_mode():
        shapes = torch.tensor([[2, 3], [1, 2]])
        mod._check_spatial_shapes(shapes, 8)
        mod._check_spatial_shapes(shapes, 8)
        with pytest.raises(AssertionError):
            mod._check_spatial_shapes(shapes, 9)
        g = _level_grids(shapes, torch.device("cpu"))
        assert _level_grids(shapes, torch.device("cpu")) is g
        mask = torch.zeros(1, 8, dtype=torch.bool)
        assert mod._no_padding(mask) and mod._no_padding(mask)
    c = PerTensorCache()
    with torch.inference_mode():
        t = torch.zeros(3)
    c.put(t, "v")
    assert c.get(t) == "v"
    m = torch.zeros(2, 4, dtype=torch.bool)
    assert mod._no_padding(m)
    m[0, 1] = True                                     # in-place change: asked again
    assert not mod._no_padding(m)


def test_flat_params_groups_clip_and_state_dict_refresh():
    """The reference's optimizer groups (src/train.py:103-117) and gradient clipping (engine.py:147-148) on the
    flat layout, against the same model trained the plain way; load_state_dict refreshes the bf16 mirror."""
    from torch import nn
    from kinet_b200.training import FlatParams, reference_param_groups

    class Net(nn.Module):
        def __init__(self):
            super().__init__()
            self.value_proj = nn.Linear(6, 6)
            self.sampling_offsets = nn.Linear(6, 4)
            self.reference_points = nn.Linear(6, 2)

        def forward(self, x):
            return self.value_proj(x).sum() + (self.sampling_offsets(x) ** 2).sum() + self.reference_points(x).sum()

    torch.manual_seed(0)
    a, b = Net(), Net()
    b.load_state_dict(a.state_dict())
    x = torch.randn(5, 6)
    ga = reference_param_groups(a, lr=1e-2, lr_linear_proj_mult=0.1)
    assert [len(g["params"]) for g in ga] == [2, 4] and ga[1]["lr"] == pytest.approx(1e-3)
    opt_a = torch.optim.AdamW(ga, lr=1e-2, weight_decay=1e-4)
    fp = FlatParams(b, groups=reference_param_groups(b, lr=1e-2, lr_linear_proj_mult=0.1))
    opt_b = torch.optim.AdamW(fp.optimizer_groups(), lr=1e-2, weight_decay=1e-4)
    for _ in range(3):
        opt_a.zero_grad()
        a(x).backward()
        torch.nn.utils.clip_grad_norm_(a.parameters(), 0.1)
        opt_a.step()
        fp.zero_grad()
        b(x).backward()
        fp.clip_grad_norm_(0.1)
        opt_b.step()
        fp.refresh()
    for (n, pa), (_, pb) in zip(a.named_parameters(), b.named_parameters()):
        assert torch.allclose(pa, pb, rtol=1e-5, atol=1e-7), n
    assert torch.allclose(fp.lp.float(), fp.master, rtol=1e-2, atol=1e-3)
    b.load_state_dict({k: torch.ones_like(v) for k, v in b.state_dict().items()})
    assert fp.lp.float().sum() == fp.master.sum() == fp.numel        # mirror follows a checkpoint load
    with pytest.raises(ValueError):
        FlatParams(Net(), groups=[{"params": list(Net().parameters())}])   # groups of another module
