"""Test infrastructure: a Deformable-DETR style encoder/decoder stack around MSDeformAttn
(BASELINE.json configs[1]: 6-layer encoder + 6-layer decoder, ~22k multi-scale tokens, 300 object queries)
with a pluggable sampling core, so the same stack can run on

  * "fused"   -- kinet_b200.MSDeformAttn with the fused prologue (the default product path),
  * "plain"   -- kinet_b200.MSDeformAttn with the reference's PyTorch prologue + MSDeformAttnFunction,
  * "refcuda" -- the reference's PyTorch prologue + the REFERENCE's own CUDA kernels compiled for sm_100a
                 (oracle/_ref/libmsda_ref_cuda.so; fp32 only, the reference has no AMP),
  * "refbwd"  -- this library's forward kernel with the reference's backward kernels.

The layer structure follows the reference's callers of the op
(src/trackformer/models/deformable_transformer.py:260-299 encoder layer, :340-386 decoder layer:
self-attention -> MSDeformAttn cross-attention -> FFN, post-norm, dropout disabled here).
It is a harness for parity and timing, not a re-implementation of the reference model.
"""
import torch
from torch import nn
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from kinet_b200.modules import MSDeformAttn
from kinet_b200.workloads import encoder_reference_points


class RefCudaFunction(Function):
    """The reference's MSDeformAttnFunction (ops/functions/ms_deform_attn_func.py:14-31) on its own kernels."""

    @staticmethod
    def forward(ctx, value, shapes, loc, attn, im2col_step):
        from oracle import msda_ref_cuda as ref
        value, loc, attn = value.float().contiguous(), loc.float().contiguous(), attn.float().contiguous()
        ctx.save_for_backward(value, shapes, loc, attn)
        return ref.ref_cuda_forward(value, shapes, loc, attn)

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_output):
        from oracle import msda_ref_cuda as ref
        value, shapes, loc, attn = ctx.saved_tensors
        gv, gl, ga = ref.ref_cuda_backward(value, shapes, loc, attn, grad_output.float().contiguous())
        return gv, None, gl, ga, None


class RefBackwardFunction(Function):
    """This library's forward, the reference kernels' backward: isolates the backward comparison from the
    (tiny, but chaotically amplified -- see test_transformer_gpu.py) forward rounding differences."""

    @staticmethod
    def forward(ctx, value, shapes, loc, attn, im2col_step):
        from kinet_b200.functions import ms_deform_attn_forward
        value, loc, attn = value.float().contiguous(), loc.float().contiguous(), attn.float().contiguous()
        ctx.save_for_backward(value, shapes, loc, attn)
        return ms_deform_attn_forward(value, shapes, loc, attn, im2col_step)

    backward = RefCudaFunction.backward


class SwitchableMSDeformAttn(MSDeformAttn):
    core = "fused"

    def forward(self, query, reference_points, input_flatten, input_spatial_shapes,
                input_padding_mask=None, query_attn_mask=None):
        if self.core in ("refcuda", "refbwd"):
            import kinet_b200.modules.ms_deform_attn as mod
            saved = mod.MSDeformAttnFunction
            mod.MSDeformAttnFunction = RefCudaFunction if self.core == "refcuda" else RefBackwardFunction
            self.fused_prologue = False
            try:
                return super().forward(query, reference_points, input_flatten, input_spatial_shapes,
                                       input_padding_mask, query_attn_mask)
            finally:
                mod.MSDeformAttnFunction = saved
        self.fused_prologue = self.core == "fused"
        return super().forward(query, reference_points, input_flatten, input_spatial_shapes,
                               input_padding_mask, query_attn_mask)


class EncoderLayer(nn.Module):
    def __init__(self, d=256, ffn=1024, L=4, M=8, P=4):
        super().__init__()
        self.self_attn = SwitchableMSDeformAttn(d, L, M, P)
        self.norm1, self.norm2 = nn.LayerNorm(d), nn.LayerNorm(d)
        self.linear1, self.linear2 = nn.Linear(d, ffn), nn.Linear(ffn, d)

    def forward(self, src, pos, ref, shapes):
        src = self.norm1(src + self.self_attn(src + pos, ref, src, shapes))
        return self.norm2(src + self.linear2(torch.relu(self.linear1(src))))


class DecoderLayer(nn.Module):
    def __init__(self, d=256, ffn=1024, L=4, M=8, P=4):
        super().__init__()
        self.cross_attn = SwitchableMSDeformAttn(d, L, M, P)
        self.self_attn = nn.MultiheadAttention(d, M, batch_first=True)
        self.norm1, self.norm2, self.norm3 = nn.LayerNorm(d), nn.LayerNorm(d), nn.LayerNorm(d)
        self.linear1, self.linear2 = nn.Linear(d, ffn), nn.Linear(ffn, d)

    def forward(self, tgt, qpos, ref, memory, shapes):
        q = tgt + qpos
        tgt = self.norm2(tgt + self.self_attn(q, q, tgt, need_weights=False)[0])
        tgt = self.norm1(tgt + self.cross_attn(tgt + qpos, ref, memory, shapes))
        return self.norm3(tgt + self.linear2(torch.relu(self.linear1(tgt))))


class Stack(nn.Module):
    def __init__(self, levels, n_enc=6, n_dec=6, n_queries=300, d=256):
        super().__init__()
        self.levels = levels
        L = len(levels)
        self.enc = nn.ModuleList(EncoderLayer(d, 1024, L) for _ in range(n_enc))
        self.dec = nn.ModuleList(DecoderLayer(d, 1024, L) for _ in range(n_dec))
        self.query_embed = nn.Embedding(n_queries, 2 * d)
        self.ref_proj = nn.Linear(d, 2)
        g = torch.Generator().manual_seed(3)
        for m in self.modules():      # non-degenerate offsets / attention (the module initialises them to 0)
            if isinstance(m, MSDeformAttn):
                with torch.no_grad():
                    m.sampling_offsets.weight.copy_(torch.randn(m.sampling_offsets.weight.shape, generator=g) * 0.01)
                    m.attention_weights.weight.copy_(torch.randn(m.attention_weights.weight.shape, generator=g) * 0.05)

    def set_core(self, core):
        for m in self.modules():
            if isinstance(m, SwitchableMSDeformAttn):
                m.core = core

    def _geometry(self, N, device):
        # built once per (batch, device): constant inputs must not be re-created (and copied from the
        # host) inside a CUDA-graph capture, and the modules validate a shapes tensor once per object
        key = (N, str(device))
        if getattr(self, "_geo_key", None) != key:
            self._geo = (torch.as_tensor(self.levels, dtype=torch.long, device=device),
                         encoder_reference_points(self.levels, N, device).contiguous())
            self._geo_key = key
        return self._geo

    def forward(self, src, pos):
        N, S, d = src.shape
        shapes, ref = self._geometry(N, src.device)
        mem = src
        for layer in self.enc:
            mem = layer(mem, pos, ref, shapes)
        qpos, tgt = self.query_embed.weight.split(d, dim=1)
        qpos, tgt = qpos[None].expand(N, -1, -1), tgt[None].expand(N, -1, -1)
        dref = self.ref_proj(qpos).sigmoid()[:, :, None, :].expand(-1, -1, len(self.levels), -1)
        for layer in self.dec:
            tgt = layer(tgt, qpos, dref, mem, shapes)
        return tgt, mem


class FusedStack(nn.Module):
    """The same 6+6-layer stack built from kinet_b200.modules' encoder / decoder classes (fused residual +
    LayerNorm + positional-add kernels between the GEMMs).  Parameter-compatible with Stack (load_from_stack)."""

    def __init__(self, levels, n_enc=6, n_dec=6, n_queries=300, d=256):
        super().__init__()
        from kinet_b200.modules import (DeformableTransformerDecoder, DeformableTransformerDecoderLayer,
                                        DeformableTransformerEncoder, DeformableTransformerEncoderLayer)
        self.levels = levels
        L = len(levels)
        self.encoder = DeformableTransformerEncoder(DeformableTransformerEncoderLayer(d, 1024, 0.0, "relu", L, 8, 4), n_enc)
        self.decoder = DeformableTransformerDecoder(DeformableTransformerDecoderLayer(d, 1024, 0.0, "relu", L, 8, 4), n_dec)
        self.query_embed = nn.Embedding(n_queries, 2 * d)
        self.ref_proj = nn.Linear(d, 2)
        self.load_from_stack(Stack(levels, n_enc, n_dec, n_queries, d))

    def load_from_stack(self, stack):
        sd = {}
        for k, v in stack.state_dict().items():
            k = k.replace("enc.", "encoder.layers.", 1) if k.startswith("enc.") else k
            k = k.replace("dec.", "decoder.layers.", 1) if k.startswith("dec.") else k
            sd[k] = v
        self.load_state_dict(sd, strict=True)

    def _geometry(self, N, device):
        key = (N, str(device))
        if getattr(self, "_geo_key", None) != key:
            self._geo = (torch.as_tensor(self.levels, dtype=torch.long, device=device),
                         torch.ones(N, len(self.levels), 2, device=device))
            self._geo_key = key
        return self._geo

    def forward(self, src, pos):
        N, S, d = src.shape
        shapes, valid = self._geometry(N, src.device)
        mem = self.encoder(src, shapes, valid, pos, None, return_stream=True)
        qpos, tgt = self.query_embed.weight.split(d, dim=1)
        qpos, tgt = qpos[None].expand(N, -1, -1), tgt[None].expand(N, -1, -1)
        dref = self.ref_proj(qpos).sigmoid()
        out, _ = self.decoder(tgt, dref, mem, shapes, valid, qpos)
        return out, mem.x


def randomise_by_name(model):
    """Parameters as a pure function of their names -- the rule of tests/golden/make_golden_transformer.py
    (case c: the two-stage model is too large to store in the golden file)."""
    import zlib
    with torch.no_grad():
        for name, p in model.named_parameters():
            gn = torch.Generator().manual_seed(zlib.crc32(name.encode()))
            if "norm" in name and name.endswith("weight"):
                p.copy_(1 + 0.1 * torch.randn(p.shape, generator=gn))
            elif p.dim() == 1:
                p.copy_(0.1 * torch.randn(p.shape, generator=gn))
            else:
                p.copy_(torch.randn(p.shape, generator=gn) * (0.5 / p.shape[-1] ** 0.5))


def two_stage_model():
    """The model of golden case c: two-stage DeformableTransformer with Linear stand-ins for the detection heads."""
    from kinet_b200.modules import DeformableTransformer
    model = DeformableTransformer(256, 8, 1, 2, 32, 0.0, "relu", True, 2, 2, 2, two_stage=True, two_stage_num_proposals=6)
    model.decoder.class_embed = nn.ModuleList(nn.Linear(256, 3) for _ in range(3))
    model.decoder.bbox_embed = nn.ModuleList(nn.Linear(256, 4) for _ in range(3))
    randomise_by_name(model)
    return model.eval()
