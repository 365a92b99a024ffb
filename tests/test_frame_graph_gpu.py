"""FrameGraphs (kinet_b200/frame_graph.py): per-frame CUDA graphs with bucketed, mask-padded query counts must
return what the eager, unpadded encoder + decoder return for every query count."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _model(dev, d=128, L=2, return_intermediate=False):
    from kinet_b200.modules import (DeformableTransformerDecoder, DeformableTransformerDecoderLayer,
                                    DeformableTransformerEncoder, DeformableTransformerEncoderLayer)
    torch.manual_seed(5)
    enc = DeformableTransformerEncoder(DeformableTransformerEncoderLayer(d, 64, 0.0, "relu", L, 8, 2), 2)
    dec = DeformableTransformerDecoder(DeformableTransformerDecoderLayer(d, 64, 0.0, "relu", L, 8, 2), 2,
                                       return_intermediate=return_intermediate)
    g = torch.Generator().manual_seed(6)
    with torch.no_grad():
        for name, p in list(enc.named_parameters()) + list(dec.named_parameters()):
            if "sampling_offsets.weight" in name or "attention_weights.weight" in name:
                p.copy_(torch.randn(p.shape, generator=g) * 0.05)
    return enc.to(dev).eval(), dec.to(dev).eval()


def _frame(dev, Lq, seed, d=128, N=1, levels=((12, 16), (6, 8)), box_refs=False):
    g = torch.Generator(device=dev).manual_seed(seed)
    S = sum(h * w for h, w in levels)
    r = lambda *s: torch.randn(*s, device=dev, generator=g)
    ref = torch.rand(N, Lq, 4 if box_refs else 2, device=dev, generator=g) * 0.6 + 0.2
    return dict(src=r(N, S, d), pos=0.3 * r(N, S, d), valid=torch.ones(N, len(levels), 2, device=dev),
                tgt=r(N, Lq, d), ref=ref, qpos=0.3 * r(N, Lq, d))


@pytest.mark.parametrize("amp", [None, torch.bfloat16])
def test_bucketed_graphs_match_eager(amp):
    from kinet_b200.frame_graph import FrameGraphs
    dev = torch.device("cuda", 0)
    enc, dec = _model(dev)
    shapes = torch.as_tensor([(12, 16), (6, 8)], dtype=torch.long, device=dev)
    fg = FrameGraphs(enc, dec, shapes, bucket=32, autocast_dtype=amp)
    tol = 1e-4 if amp is None else 3e-2
    for k, Lq in enumerate([5, 31, 32, 33, 70, 7, 64]):
        f = _frame(dev, Lq, seed=k)
        hs, ref_out, mem = fg(f["src"], f["pos"], f["valid"], f["tgt"], f["ref"], f["qpos"])
        with torch.no_grad(), torch.autocast("cuda", dtype=amp or torch.bfloat16, enabled=amp is not None):
            mem0 = enc(f["src"], shapes, f["valid"], f["pos"], None)
            hs0, ref0 = dec(f["tgt"], f["ref"], mem0, shapes, f["valid"], f["qpos"])
        assert hs.shape == hs0.shape == (1, Lq, 128)
        assert float((mem - mem0).abs().max() / mem0.abs().max()) < tol
        assert float((hs - hs0).abs().max() / hs0.abs().max()) < tol, Lq
        assert torch.equal(ref_out, ref0)
    assert fg.recorded_buckets == [32, 64, 96] and fg.replays == 7


def test_masks_intermediates_and_box_references():
    from kinet_b200.frame_graph import FrameGraphs
    dev = torch.device("cuda", 0)
    enc, dec = _model(dev, return_intermediate=True)
    shapes = torch.as_tensor([(12, 16), (6, 8)], dtype=torch.long, device=dev)
    fg = FrameGraphs(enc, dec, shapes, bucket=16, autocast_dtype=None)
    for k, Lq in enumerate([9, 20]):
        f = _frame(dev, Lq, seed=10 + k, N=2, box_refs=True)
        S = f["src"].shape[1]
        pmask = torch.zeros(2, S, dtype=torch.bool, device=dev)
        pmask[1, -5:] = True
        qmask = torch.zeros(2, Lq, dtype=torch.bool, device=dev)
        qmask[0, -2:] = True
        hs, ref_out, mem = fg(f["src"], f["pos"], f["valid"], f["tgt"], f["ref"], f["qpos"], pmask, qmask)
        with torch.no_grad():
            mem0 = enc(f["src"], shapes, f["valid"], f["pos"], pmask)
            hs0, ref0 = dec(f["tgt"], f["ref"], mem0, shapes, f["valid"], f["qpos"], pmask, qmask)
        assert hs.shape == hs0.shape == (2, 2, Lq, 128)
        keep = ~qmask                                    # masked queries' own rows are not compared
        assert float((hs[:, keep] - hs0[:, keep]).abs().max() / hs0[:, keep].abs().max()) < 1e-4
        assert float((mem - mem0).abs().max()) < 1e-4 * float(mem0.abs().max())
        assert torch.equal(ref_out, ref0)


def test_cached_low_precision_weights_follow_updates():
    from kinet_b200.functions.layer_ops import linear
    dev = torch.device("cuda", 0)
    lin = torch.nn.Linear(64, 32).to(dev)
    x = torch.randn(5, 64, device=dev)
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        a = linear(x, lin)
        assert a.dtype == torch.bfloat16 and torch.equal(a, lin(x))
        cached = lin.__dict__["_kinet_lp"][1]
        assert linear(x, lin) is not None and lin.__dict__["_kinet_lp"][1] is cached      # reused
        lin.weight.mul_(2.0)                                                               # in-place update
        b = linear(x, lin)
        # (autocast's own cache still holds the OLD weight copy inside this region: compare with an explicit cast)
        ref = torch.nn.functional.linear(x.bfloat16(), lin.weight.bfloat16(), lin.bias.bfloat16())
        assert lin.__dict__["_kinet_lp"][1] is not cached and torch.equal(b, ref)
