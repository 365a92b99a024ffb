"""Synthetic workloads for tests and bench.py: the pyramids and shapes of SURVEY.md section 8
(BASELINE.json configs) and input generators for the two sampling regimes.

Pyramids follow the reference's resize rule and strides 8/16/32/64
(datasets/transforms.py:118-136, deformable_detr.py:55-72).
"""
import math

import torch

PYRAMIDS = {
    # 800x1333 (config 1/2/4): Deformable-DETR default test size
    "detr_800x1333": [(100, 167), (50, 84), (25, 42), (13, 21)],
    # MOT17 1080x1920 -> 750x1333 (config 3)
    "mot17_750x1333": [(94, 167), (47, 84), (24, 42), (12, 21)],
    # MOT20 full-res 1080x1920 (config 5)
    "mot20_1080x1920": [(135, 240), (68, 120), (34, 60), (17, 30)],
}

# tag -> dict(pyramid, frames (levels repeated), N, M, D, P, Lq ('S' = encoder form))
SHAPES = {
    "C1_dec": dict(pyramid="detr_800x1333", frames=1, N=2, M=8, D=32, P=4, Lq=300),
    "C2_enc": dict(pyramid="detr_800x1333", frames=1, N=2, M=8, D=32, P=4, Lq="S"),
    "C2_dec": dict(pyramid="detr_800x1333", frames=1, N=2, M=8, D=32, P=4, Lq=300),
    "C3_enc": dict(pyramid="mot17_750x1333", frames=1, N=1, M=8, D=36, P=4, Lq="S"),
    "C3_dec": dict(pyramid="mot17_750x1333", frames=2, N=1, M=8, D=36, P=4, Lq=560),
    "C5_enc": dict(pyramid="mot20_1080x1920", frames=1, N=1, M=8, D=32, P=4, Lq="S"),
    "C5_dec": dict(pyramid="mot20_1080x1920", frames=1, N=1, M=8, D=32, P=4, Lq=800),
}


def resolve(tag):
    s = dict(SHAPES[tag])
    levels = PYRAMIDS[s["pyramid"]] * s["frames"]
    s["levels"] = levels
    s["L"] = len(levels)
    s["S"] = sum(h * w for h, w in levels)
    s["encoder"] = s["Lq"] == "S"
    if s["encoder"]:
        s["Lq"] = s["S"]
    return s


def algorithmic_bytes(s, e_value, e_aux):
    """compulsory bytes of one forward and one backward call."""
    N, S, M, D, L, P, Lq = s["N"], s["S"], s["M"], s["D"], s["L"], s["P"], s["Lq"]
    samples = N * Lq * M * L * P
    val = e_value * min(N * S * M * D, 4 * samples * D)
    loc = e_aux * 2 * samples
    att = e_aux * samples
    out = e_value * N * Lq * M * D
    fwd = val + loc + att + out
    bwd = (val + loc + att) + out + 2 * val + loc + att
    return fwd, bwd


def encoder_reference_points(levels, N, device, dtype=torch.float32):
    """Pixel-centre grid of every level, broadcast to all levels (valid_ratio = 1), like
    DeformableTransformerEncoder.get_reference_points (deformable_transformer.py:308-321)."""
    pts = []
    for h, w in levels:
        ys = (torch.arange(h, device=device, dtype=dtype) + 0.5) / h
        xs = (torch.arange(w, device=device, dtype=dtype) + 0.5) / w
        yy, xx = torch.meshgrid(ys, xs, indexing="ij")
        pts.append(torch.stack([xx.reshape(-1), yy.reshape(-1)], -1))
    ref = torch.cat(pts, 0)                       # (S, 2)
    return ref[None, :, None, :].expand(N, -1, len(levels), -1)


def make_inputs(tag_or_shape, device, value_dtype=torch.float32, aux_dtype=torch.float32,
                regime="structured", seed=0):
    """-> value (N,S,M,D), shapes (L,2) int64, loc (N,Lq,M,L,P,2), attn (N,Lq,M,L,P).

    regime 'structured': what the model produces at initialisation -- reference point (pixel grid
    for the encoder form, U(0.05,0.95) for the decoder form) plus the module's compass-direction
    offsets (p+1 pixels, ms_deform_attn.py:36-41) plus N(0, 1 px) noise.
    regime 'uniform': loc ~ U(-0.05, 1.05): worst-case locality, exercises zero padding.
    """
    s = resolve(tag_or_shape) if isinstance(tag_or_shape, str) else tag_or_shape
    N, S, M, D, L, P, Lq = s["N"], s["S"], s["M"], s["D"], s["L"], s["P"], s["Lq"]
    levels = s["levels"]
    g = torch.Generator(device=device).manual_seed(seed)
    value = torch.randn(N, S, M, D, generator=g, device=device, dtype=torch.float32).to(value_dtype)
    shapes = torch.as_tensor(levels, dtype=torch.int64, device=device)
    attn = torch.softmax(torch.randn(N, Lq, M, L * P, generator=g, device=device), -1)
    attn = attn.view(N, Lq, M, L, P).to(aux_dtype)
    if regime == "uniform":
        loc = torch.rand(N, Lq, M, L, P, 2, generator=g, device=device) * 1.1 - 0.05
    else:
        if s["encoder"]:
            ref = encoder_reference_points(levels, N, device)          # (N, S, L, 2)
        else:
            ref = torch.rand(N, Lq, 1, 2, generator=g, device=device) * 0.9 + 0.05
            ref = ref.expand(-1, -1, L, -1)
        th = torch.arange(M, device=device, dtype=torch.float32) * (2.0 * math.pi / M)
        dirs = torch.stack([th.cos(), th.sin()], -1)
        dirs = dirs / dirs.abs().max(-1, keepdim=True)[0]              # compass directions
        scale = torch.arange(1, P + 1, device=device, dtype=torch.float32)
        off = dirs.view(1, 1, M, 1, 1, 2) * scale.view(1, 1, 1, 1, P, 1)
        off = off + torch.randn(N, Lq, M, L, P, 2, generator=g, device=device)
        wh = torch.as_tensor([[w, h] for h, w in levels], device=device, dtype=torch.float32)
        loc = ref[:, :, None, :, None, :] + off / wh.view(1, 1, 1, L, 1, 2)
    return value, shapes, loc.to(aux_dtype).contiguous(), attn.contiguous()
