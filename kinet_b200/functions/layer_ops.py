"""Autograd wrappers over the layer glue kernels (include/kinet_layers_b200.h, csrc/layer_ops.cu).

``add_layer_norm`` fuses what the reference's encoder / decoder layers do between two GEMMs
(src/trackformer/models/deformable_transformer.py:294-296, 282-284, 363-365, 375-376, 382-383):
``x = norm(x + dropout(branch))`` followed by ``with_pos_embed(x, pos)`` for the next attention's
query (:277-278, 293, 369, 379), plus the dtype casts autocast would put in front of the next Linear
layers.  ``linear`` is ``F.linear`` whose bias gradient is a fixed-order fp32 column sum instead of
PyTorch's generic reduce kernel.  CUDA only -- CPU tensors raise, like the op itself.
"""
import ctypes

import torch
import torch.nn.functional as F
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from .. import _lib

_LN_DTYPES = (torch.float32, torch.bfloat16, torch.float16)
MAX_C = 1024


def _ptr(t):
    return None if t is None else t.data_ptr()


def _code(t, default=0):
    return default if t is None else _lib.dtype_code(t.dtype)


def _rows(t, C):
    if t.shape[-1] != C or not t.is_contiguous():
        raise RuntimeError("kinet_b200: expected a contiguous tensor with %d channels last" % C)
    return t.numel() // C


def _canon_pos(pos, x):
    """pos as a contiguous [pos_rows, C] block that row i of x reads at i % pos_rows.
    Accepts pos shaped like x, or broadcast over the leading (batch) dimension."""
    if pos.shape[-1] != x.shape[-1]:
        raise RuntimeError("kinet_b200: pos must have the channel count of x")
    if pos.dim() == x.dim() and pos.shape != x.shape and pos.shape[0] == 1:
        pos = pos[0]
    if pos.dim() == x.dim() and pos.shape == x.shape and x.dim() >= 2 and pos.stride(0) == 0:
        pos = pos[0]                                  # an .expand()ed batch dimension
    if pos.dim() == x.dim():
        if pos.shape != x.shape:
            raise RuntimeError("kinet_b200: pos %s does not match x %s" % (tuple(pos.shape), tuple(x.shape)))
    elif tuple(pos.shape) != tuple(x.shape[x.dim() - pos.dim():]):
        raise RuntimeError("kinet_b200: pos %s does not broadcast over x %s" % (tuple(pos.shape), tuple(x.shape)))
    return pos.contiguous()


# ---- dropout inside the add + LayerNorm kernels ---------------------------------------------------------------------
# The kernels decide which elements of the branch survive from a hash of (element index, seed, site, step) and the
# backward regenerates the mask (include/kinet_layers_b200.h).  seed: per process (from torch.initial_seed(), i.e.
# torch.manual_seed controls it); site: a running number per forward call, so eager calls never share a mask; step: a
# DEVICE counter that a CUDA-graph training step advances inside the graph (advance_dropout_step), so every replay
# draws new masks although the sites are baked into the graph.
_drop_state = {"seed": None, "site": 0, "steps": {}}


def set_dropout_seed(seed):
    _drop_state["seed"] = int(seed) & 0xFFFFFFFFFFFFFFFF


def _dropout_seed():
    if _drop_state["seed"] is None:
        set_dropout_seed(torch.initial_seed() * 0x9E3779B97F4A7C15 + 0x1234567)
    return _drop_state["seed"]


def dropout_step_tensor(device):
    """The int64 device counter the kernels read (one per device)."""
    key = str(device)
    t = _drop_state["steps"].get(key)
    if t is None:
        t = _drop_state["steps"][key] = torch.zeros(1, dtype=torch.int64, device=device)
    return t


def advance_dropout_step(device):
    """New masks from here on (an in-place add on the device: capturable, replays advance it too)."""
    dropout_step_tensor(device).add_(1)


def _next_site():
    _drop_state["site"] = (_drop_state["site"] + 1) & 0xFFFFFFFF
    return _drop_state["site"]


class AddLayerNormFunction(Function):
    """(y, y_lp, ypos_lp) = AddLayerNormFunction.apply(x, r, pos, gamma, beta, eps, lp_dtype, want_lp,
                                                       drop_p=0.0, branch_bias=None)

    y = LayerNorm(x + dropout(r, drop_p)) in x.dtype; y_lp = y.to(lp_dtype) if want_lp else None;
    ypos_lp = (y + pos).to(lp_dtype) if pos is not None else None.  r may be None.
    branch_bias: the bias parameter of the Linear layer that produced r (already added by that layer: it is NOT
    added here); its gradient -- the column sums of d r -- comes out of this function's backward, so that layer can
    skip its own bias-gradient pass (linear(..., skip_bias_grad=True))."""

    @staticmethod
    def forward(ctx, x, r, pos, gamma, beta, eps, lp_dtype, want_lp, drop_p=0.0, branch_bias=None):
        if not x.is_cuda:
            raise RuntimeError("Not implemented on the CPU")
        C = x.shape[-1]
        x = x.contiguous()
        rows = _rows(x, C)
        if r is not None:
            r = r.contiguous()
            if r.shape != x.shape:
                raise RuntimeError("kinet_b200: branch output %s does not match x %s" % (tuple(r.shape), tuple(x.shape)))
        gamma32, beta32 = gamma.detach().float().contiguous(), beta.detach().float().contiguous()
        posc = _canon_pos(pos.detach(), x) if pos is not None else None
        need_stats = any(ctx.needs_input_grad[:5])
        y = torch.empty_like(x)
        y_lp = torch.empty(x.shape, dtype=lp_dtype, device=x.device) if want_lp else None
        ypos = torch.empty(x.shape, dtype=lp_dtype, device=x.device) if posc is not None else None
        need_stats = need_stats or (branch_bias is not None and ctx.needs_input_grad[9])
        stats = torch.empty(rows, 2, dtype=torch.float32, device=x.device) if need_stats else None
        drop_p = float(drop_p) if r is not None else 0.0
        drop = (0.0, 0, 0, None)
        if drop_p > 0.0:
            drop = (drop_p, _next_site(), _dropout_seed(), dropout_step_tensor(x.device))
        a = _lib.KinetLnFwdArgs(
            _ptr(x), _code(x), _ptr(r), _code(r), _ptr(posc), _code(posc),
            (posc.numel() // C) if posc is not None else 1,
            _ptr(gamma32), _ptr(beta32), _ptr(y), _code(y), _ptr(y_lp), _ptr(ypos),
            _lib.dtype_code(lp_dtype), _ptr(stats), rows, C, float(eps),
            drop[0], drop[1], drop[2], _ptr(drop[3]))
        lib = _lib.load()
        with _lib.device_guard(x.device):
            rc = lib.kinet_b200_add_layernorm_forward(ctypes.byref(a), _lib.raw_stream(x.device))
        _lib.check(rc, "kinet_b200_add_layernorm_forward")
        ctx.set_materialize_grads(False)
        if need_stats:
            ctx.save_for_backward(x, r, gamma32, stats)
            ctx.pos_meta = None if pos is None else (pos.shape, pos.dtype, posc.shape)
            ctx.param_dtypes = (gamma.dtype, beta.dtype)
            ctx.drop = drop
            ctx.bias_dtype = None if branch_bias is None or r is None else branch_bias.dtype
        return y, y_lp, ypos

    @staticmethod
    @once_differentiable
    def backward(ctx, dy, dy_lp, dypos):
        x, r, gamma32, stats = ctx.saved_tensors
        if dy is None and dy_lp is None and dypos is None:
            return (None,) * 10
        C = x.shape[-1]
        rows = x.numel() // C
        lp = [t for t in (dy_lp, dypos) if t is not None]
        if len(lp) == 2 and lp[0].dtype != lp[1].dtype:
            raise RuntimeError("kinet_b200: companion gradients must share a dtype")
        dy = dy.contiguous() if dy is not None else None
        dy_lp = dy_lp.contiguous() if dy_lp is not None else None
        dypos = dypos.contiguous() if dypos is not None else None
        dx = torch.empty_like(x)
        dr = torch.empty_like(r) if (r is not None and ctx.needs_input_grad[1]) else None
        want_dbias = ctx.bias_dtype is not None and ctx.needs_input_grad[9]
        ncol = 3 if want_dbias else 2
        dgb = torch.empty(ncol, C, dtype=torch.float32, device=x.device)
        lib = _lib.load()
        partial = torch.empty(lib.kinet_b200_colsum_partial_floats(ncol * C), dtype=torch.float32, device=x.device)
        drop = ctx.drop
        a = _lib.KinetLnBwdArgs(
            _ptr(dy), _code(dy), _ptr(dy_lp), _ptr(dypos), _code(lp[0]) if lp else 0,
            _ptr(x), _code(x), _ptr(r), _code(r), _ptr(gamma32), _ptr(stats),
            _ptr(dx), _ptr(dr), _ptr(dgb), _ptr(partial), rows, C,
            1 if want_dbias else 0, drop[0], drop[1], drop[2], _ptr(drop[3]))
        with _lib.device_guard(x.device):
            rc = lib.kinet_b200_add_layernorm_backward(ctypes.byref(a), _lib.raw_stream(x.device))
        _lib.check(rc, "kinet_b200_add_layernorm_backward")
        dpos = None
        if ctx.needs_input_grad[2] and dypos is not None:
            shape, dtype, cshape = ctx.pos_meta
            if tuple(shape) == tuple(x.shape):      # incl. an expanded batch dimension: autograd sums it
                dpos = dypos.to(dtype)
            else:
                dpos = dypos.reshape((-1,) + tuple(cshape)).sum(0).to(dtype).reshape(shape)
        gd, bd = ctx.param_dtypes
        return (dx if ctx.needs_input_grad[0] else None, dr, dpos,
                dgb[0].to(gd) if ctx.needs_input_grad[3] else None,
                dgb[1].to(bd) if ctx.needs_input_grad[4] else None, None, None, None, None,
                dgb[2].to(ctx.bias_dtype) if want_dbias else None)


def add_layer_norm(x, r, norm, pos=None, lp_dtype=None, want_lp=False, drop_p=0.0, branch_bias=None):
    """norm(x + dropout(r, drop_p)) with the layer's nn.LayerNorm parameters -> (y, y_lp, ypos_lp); see
    AddLayerNormFunction."""
    if tuple(norm.normalized_shape) != (x.shape[-1],) or norm.weight is None or norm.bias is None:
        raise RuntimeError("kinet_b200: add_layer_norm needs an affine LayerNorm over the last dimension")
    return AddLayerNormFunction.apply(x, r, pos, norm.weight, norm.bias, norm.eps,
                                      lp_dtype if lp_dtype is not None else x.dtype, want_lp, drop_p, branch_bias)


def column_sum(a2d, out_dtype=None):
    """Fixed-order fp32 column sums of a contiguous (rows, C) CUDA matrix."""
    if not a2d.is_cuda:
        raise RuntimeError("Not implemented on the CPU")
    rows, C = a2d.shape
    out_dtype = out_dtype or a2d.dtype
    lib = _lib.load()
    out = torch.empty(C, dtype=out_dtype, device=a2d.device)
    partial = torch.empty(lib.kinet_b200_colsum_partial_floats(C), dtype=torch.float32, device=a2d.device)
    with _lib.device_guard(a2d.device):
        rc = lib.kinet_b200_colsum(_ptr(a2d), _lib.dtype_code(a2d.dtype), rows, C, _ptr(out),
                                   _lib.dtype_code(out_dtype), _ptr(partial), _lib.raw_stream(a2d.device))
    _lib.check(rc, "kinet_b200_colsum")
    return out


def colsum_supported(a2d):
    return (a2d.is_cuda and a2d.dtype in _LN_DTYPES and a2d.dim() == 2 and a2d.is_contiguous()
            and a2d.shape[1] % 4 == 0 and 0 < a2d.shape[1] <= MAX_C)


def relu_backward_column_sum(dy2d, h2d, out_dtype, drop_p=0.0):
    """dh = dy * (h > 0) [/ (1 - p') when h = dropout(relu(a), drop_p), see dropout_] and its column sums in one pass
    (h = the ReLU's output) -> (dh, sums)."""
    rows, C = dy2d.shape
    lib = _lib.load()
    dh = torch.empty_like(dy2d)
    out = torch.empty(C, dtype=out_dtype, device=dy2d.device)
    partial = torch.empty(lib.kinet_b200_colsum_partial_floats(C), dtype=torch.float32, device=dy2d.device)
    with _lib.device_guard(dy2d.device):
        rc = lib.kinet_b200_relu_dropout_backward_colsum(_ptr(dy2d), _ptr(h2d), float(drop_p), _lib.dtype_code(dy2d.dtype),
                                                         rows, C, _ptr(dh), _ptr(out), _lib.dtype_code(out_dtype),
                                                         _ptr(partial), _lib.raw_stream(dy2d.device))
    _lib.check(rc, "kinet_b200_relu_dropout_backward_colsum")
    return dh, out


def dropout_(x, p):
    """In-place dropout of a contiguous CUDA tensor without a mask tensor (kinet_b200_dropout_forward): the keep / drop
    decisions are a hash of (element index, seed, call site, device step counter), like the add + LayerNorm kernels'."""
    if p <= 0.0 or x.numel() == 0:
        return x
    if not (x.is_cuda and x.is_contiguous() and x.dtype in _LN_DTYPES and x.numel() % 4 == 0):
        raise RuntimeError("kinet_b200: dropout_ needs a contiguous fp32 / bf16 / fp16 CUDA tensor of 4 n elements")
    lib = _lib.load()
    with _lib.device_guard(x.device):
        rc = lib.kinet_b200_dropout_forward(_ptr(x), _ptr(x), _lib.dtype_code(x.dtype), x.numel(), float(p), _next_site(),
                                            _dropout_seed(), _ptr(dropout_step_tensor(x.device)), _lib.raw_stream(x.device))
    _lib.check(rc, "kinet_b200_dropout_forward")
    return x


def dropout_scale(p):
    """1 / (1 - p') with p' = p rounded to 1 / 65536, as the kernels apply it."""
    return 65536.0 / (65536 - min(int(p * 65536 + 0.5), 65535)) if p > 0 else 1.0


class LinearFunction(Function):
    """y = act(x @ weight.T + bias), act = identity or ReLU (cuBLAS GEMMs, the ReLU in the GEMM's epilogue).
    Under CUDA autocast the operands are cast to the autocast dtype like F.linear would; ``w_lp`` / ``b_lp``
    are optional ready-made low-precision copies (training.FlatParams keeps them current with one cast kernel
    per step).  Backward: two cuBLAS GEMMs (the weight gradient leaves the GEMM in the parameter's dtype) and
    ONE fused pass for the bias gradient (+ the ReLU's backward)."""

    @staticmethod
    def forward(ctx, x, weight, bias, relu, w_lp, b_lp, skip_bias_grad=False, drop_p=0.0):
        """drop_p (with relu): y = dropout(relu(x @ W.T + b), drop_p), the FFN's inner dropout, without a mask tensor."""
        ctx.skip_bias_grad = bool(skip_bias_grad)      # the bias gradient comes out of the next add + LayerNorm's backward
        lp = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else None
        w, b, x_dtype = weight, bias, x.dtype
        if lp is not None:
            x = x.to(lp)
            w, b = (w_lp, b_lp) if (w_lp is not None and w_lp.dtype == lp) else (weight.to(lp), bias.to(lp))
        elif x.dtype != weight.dtype:
            x = x.to(weight.dtype)
        x2 = x.reshape(-1, x.shape[-1])
        with torch.autocast("cuda", enabled=False):
            if relu:
                y2 = torch._addmm_activation(b, x2, w.t(), use_gelu=False)
            else:
                y2 = torch.addmm(b, x2, w.t())
        drop_p = float(drop_p) if relu else 0.0
        if drop_p > 0.0:
            dropout_(y2, drop_p)               # in place: the backward only needs dropout(relu(.)) (its sign pattern)
        ctx.save_for_backward(x2, w, y2 if relu else None)
        ctx.drop_p = drop_p
        ctx.meta = (weight.dtype, bias.dtype, x_dtype, x.shape, bool(relu))
        return y2.view(x.shape[:-1] + (w.shape[0],))

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        x2, w, y2 = ctx.saved_tensors
        wd, bd, x_dtype, x_shape, relu = ctx.meta
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        fused = colsum_supported(dy2)
        db = None
        if relu:
            if fused:
                dy2, db = relu_backward_column_sum(dy2, y2, bd, ctx.drop_p)
            else:
                dy2 = dy2 * (y2 > 0).to(dy2.dtype) * dropout_scale(ctx.drop_p)
        if ctx.skip_bias_grad:
            db = None
        elif db is None and ctx.needs_input_grad[2]:
            db = column_sum(dy2, bd) if fused else dy2.sum(0).to(bd)
        dx = (dy2 @ w).view(x_shape).to(x_dtype) if ctx.needs_input_grad[0] else None
        dw = None
        if ctx.needs_input_grad[1]:
            if wd == torch.float32 and dy2.dtype in (torch.bfloat16, torch.float16):
                dw = torch.mm(dy2.t(), x2, out_dtype=torch.float32)    # fp32 straight out of the GEMM: no cast pass
            else:
                dw = (dy2.t() @ x2).to(wd)
        return dx, dw, db if ctx.needs_input_grad[2] else None, None, None, None, None, None


def lp_params(owner, weight, bias, dtype):
    """Low-precision copies of a module's (weight, bias) for inference under autocast, made once and kept on
    ``owner`` until a parameter changes: in-place updates through the parameter bump ``_version``, re-assignment
    changes the identity, ``p.data = ...`` / ``module.to()`` / ``module.half()`` change ``data_ptr``.  In-place edits
    THROUGH ``.data`` (``p.data.copy_()``, ``constant_(p.data, ...)``; EMA weight swaps) change none of the three:
    call ``invalidate_lp_cache(module)`` after those.  Unlike autocast's own cache they outlive the autocast region (and cover parameters that are
    used through slices, which autocast never caches), so a per-frame CUDA graph recorded after one eager
    warm-up holds no weight casts.  A graph keeps reading the copies it was recorded with: re-record it after
    changing the weights."""
    key = (dtype, id(weight), weight._version, weight.data_ptr(),
           id(bias), bias._version if bias is not None else 0, bias.data_ptr() if bias is not None else 0)
    hit = owner.__dict__.get("_kinet_lp")
    if hit is None or hit[0] != key:
        hit = (key, weight.detach().to(dtype), bias.detach().to(dtype) if bias is not None else None)
        owner.__dict__["_kinet_lp"] = hit
    return hit[1], hit[2]


def invalidate_lp_cache(module):
    """Drop the cached low-precision parameter copies of ``module`` and all its submodules (after parameter edits
    that ``lp_params`` cannot see, e.g. ``p.data.copy_(...)``)."""
    for m in module.modules():
        m.__dict__.pop("_kinet_lp", None)


def inference_mirror(owner, weight, bias, x):
    """(weight, bias) copies to use for ``x`` right now: FlatParams' mirror when the module has one, the cached
    inference copies under no-grad autocast, else None (use the parameters themselves)."""
    mirror = owner.__dict__.get("_kinet_mirror")
    if mirror is None and x.is_cuda and not torch.is_grad_enabled() and torch.is_autocast_enabled("cuda") \
            and bias is not None:
        mirror = lp_params(owner, weight, bias, torch.get_autocast_dtype("cuda"))
    return mirror


def linear_params(x, weight, bias, relu=False, mirror=None, skip_bias_grad=False, drop_p=0.0):
    """act(x @ weight.T + bias) on explicit parameter tensors (e.g. slices of an in_proj_weight).
    mirror: optional (weight, bias) low-precision copies kept current by the caller.  skip_bias_grad: the caller
    obtains the bias gradient elsewhere (add_layer_norm(..., branch_bias=bias)).  drop_p: dropout behind the ReLU."""
    if x.is_cuda and torch.is_grad_enabled() and bias is not None and (weight.requires_grad or bias.requires_grad):
        w_lp, b_lp = mirror if mirror is not None else (None, None)
        return LinearFunction.apply(x, weight, bias, relu, w_lp, b_lp, skip_bias_grad and not relu, drop_p if relu else 0.0)
    if drop_p > 0.0:
        return F.dropout(linear_params(x, weight, bias, relu, mirror), drop_p, True)
    if mirror is not None and x.is_cuda and torch.is_autocast_enabled("cuda") \
            and mirror[0].dtype == torch.get_autocast_dtype("cuda"):
        weight, bias = mirror
        x = x if x.dtype == weight.dtype else x.to(weight.dtype)
    if relu and x.is_cuda and bias is not None and x.dtype == weight.dtype and not torch.is_grad_enabled():
        y2 = torch._addmm_activation(bias, x.reshape(-1, x.shape[-1]), weight.t(), use_gelu=False)
        return y2.view(x.shape[:-1] + (weight.shape[0],))
    y = F.linear(x, weight, bias)
    return F.relu(y) if relu else y


def linear(x, layer, relu=False, skip_bias_grad=False, drop_p=0.0):
    """``act(layer(x))`` for an nn.Linear, act = identity or ReLU.  With parameter gradients: LinearFunction.
    Without (inference): one GEMM (ReLU in its epilogue), under CUDA autocast on low-precision parameter copies
    cached on the layer (or kept current by training.FlatParams)."""
    return linear_params(x, layer.weight, layer.bias, relu, inference_mirror(layer, layer.weight, layer.bias, x),
                         skip_bias_grad, drop_p)
