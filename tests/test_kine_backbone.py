"""The `kine` model's token backbones and positional encoding (kinet_b200/modules/kinematic_backbone.py; reference
src/trackformer/models/backbone.py:110-212, position_encoding.py:152-185) against golden vectors made from the
reference's own classes (tests/golden/make_golden_kine_backbone.py), and the whole `kine` detector from raw detections."""
import os

import numpy as np
import pytest
import torch

from test_kine import _by_name, _rel

HERE = os.path.dirname(os.path.abspath(__file__))


def _z():
    with np.load(os.path.join(HERE, "golden", "kine_backbone_f32.npz")) as z:
        return {k: z[k] for k in z.files}


def _backbones(dev, max_det=7, d=72):
    from kinet_b200.modules.kinematic_backbone import build_kine_backbone
    back = build_kine_backbone(d, max_number_detection=max_det)
    for b in back:
        _by_name(b).to(dev).eval()
    return back


def test_backbones_and_detection_rank_embedding_match_reference_golden_on_cpu():
    """case f: Joiner(Kinet_Backbone, PositionEmbeddingSineDetections) for detections and metadata, fp32 on the CPU."""
    from kinet_b200.modules.kinematic_backbone import NestedTensor
    z = _z()
    back = _backbones(torch.device("cpu"))
    mask = torch.from_numpy(z["f:mask"])
    with torch.no_grad():
        for name, b, key in (("det", back[0], "boxes"), ("meta", back[1], "conf")):
            feats, pos = b(NestedTensor(torch.from_numpy(z["f:" + key]), mask))
            assert len(feats) == 1 and feats[0].mask is mask
            assert _rel(feats[0].tensors, torch.from_numpy(z["f:tokens_" + name])) < 2e-6, name
            assert _rel(pos[0], torch.from_numpy(z["f:pos_" + name])) < 2e-6, name
    # the rank embedding wraps at max_detections and ignores padded detections
    assert torch.equal(pos[0][0, 0], pos[0][0, 7]) and not torch.equal(pos[0][0, 0], pos[0][0, 1])
    assert torch.equal(pos[0][1, 5], pos[0][1, 8])               # the padded tail repeats the last real rank


def test_backbone_classes_mirror_the_reference():
    """Same constructor signatures and state_dict keys as the reference's classes (when /root/reference is there)."""
    import inspect
    import sys
    if not os.path.isdir("/root/reference"):
        pytest.skip("reference sources not present")
    sys.path.insert(0, os.path.join(HERE, "golden"))
    import importlib
    from make_golden_layers import load_reference
    load_reference()
    bb = importlib.import_module("trackformer.models.backbone")
    pe = importlib.import_module("trackformer.models.position_encoding")
    from kinet_b200.modules import kinematic_backbone as kb
    for name, ref in (("LayerBackboneFC", bb), ("Kinet_Backbone", bb), ("Joiner", bb), ("PositionEmbeddingSineDetections", pe)):
        a, b = getattr(kb, name), getattr(ref, name)
        assert list(inspect.signature(a.__init__).parameters) == list(inspect.signature(b.__init__).parameters), name
        assert list(inspect.signature(a.forward).parameters) == list(inspect.signature(b.forward).parameters), name
    ours = kb.Joiner(kb.Kinet_Backbone(4, [256, 512, 72]), kb.PositionEmbeddingSineDetections(36)).state_dict()
    theirs = bb.Joiner(bb.Kinet_Backbone(4, [256, 512, 72]), pe.PositionEmbeddingSineDetections(36)).state_dict()
    assert list(ours) == list(theirs) and all(ours[k].shape == theirs[k].shape for k in ours)


@pytest.mark.gpu
def test_whole_kine_model_from_raw_detections_matches_reference_golden():
    """case g: KinematicDetectorTransformer over the real backbones (boxes + confidences in, logits + boxes out), fp32;
    and the same call under bf16 autocast stays close."""
    from kinet_b200.modules import kinematic_transformer as kt
    from kinet_b200.modules.kinematic_backbone import NestedTensor, NestedTensorKinet
    from kinet_b200.modules.kinematic_detector import KinematicDetectorTransformer
    dev = torch.device("cuda:0")
    z = _z()
    t = lambda k: torch.from_numpy(z[k]).to(dev)
    transformer = kt.DualKinematicTransformer(d_model=72, nhead=8, num_encoder_layers=1, num_decoder_layers=2,
                                              dim_feedforward=48, dropout=0.0, return_intermediate_dec=True)
    model = KinematicDetectorTransformer(_backbones(dev), transformer, num_classes=1, num_queries=6, aux_loss=True,
                                         dim_tracklets_det=16, dim_tracklets_metadata=8)
    _by_name(model).to(dev).eval()
    smp = NestedTensorKinet(NestedTensor(t("f:boxes"), t("f:mask")), NestedTensor(t("f:conf"), t("f:mask")))
    with torch.no_grad():
        out, _, _, src_out, hs_det = model(smp, None)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out_lp = model(smp, None)[0]
    assert _rel(src_out, t("g:src_out")) < 5e-6
    for got, k in ((out["pred_logits"], "pred_logits"), (out["pred_boxes"], "pred_boxes"),
                   (out["aux_outputs"][0]["pred_logits"], "aux_logits"), (hs_det, "hs_det")):
        assert _rel(got, t("g:" + k)) < 2e-5, k
    assert _rel(out_lp["pred_boxes"], t("g:pred_boxes")) < 3e-2 and _rel(out_lp["pred_logits"], t("g:pred_logits")) < 6e-2


@pytest.mark.gpu
def test_kine_tracker_from_raw_detections_graphs_match_eager():
    """KineSequenceTracker fed the frames' RAW detections (boxes + confidences as host arrays): with graphs the frame is one
    pinned block + one replay (backbones, masks, trail projections, transformer, heads recorded together); same track ids
    and boxes as the eager loop through KinematicDetectorTransformer.forward(NestedTensorKinet, targets)."""
    from kinet_b200.modules import kinematic_transformer as kt
    from kinet_b200.modules.kinematic_detector import KinematicDetectorTransformer
    from kinet_b200.tracking import KineSequenceTracker
    dev = torch.device("cuda:0")
    d, nq, n_frames = 72, 12, 5
    transformer = kt.DualKinematicTransformer(d_model=d, nhead=8, num_encoder_layers=1, num_decoder_layers=1,
                                              dim_feedforward=48, dropout=0.0, return_intermediate_dec=True)
    det = _by_name(KinematicDetectorTransformer(_backbones(dev, max_det=60), transformer, num_classes=1, num_queries=nq,
                                                dim_tracklets_det=4 * n_frames, dim_tracklets_metadata=n_frames)).to(dev).eval()
    with torch.no_grad():
        det.class_embed.weight.mul_(6.0)
        det.bbox_embed.layers[-1].weight.mul_(8.0)
    rng = np.random.default_rng(5)
    scene = rng.random((20, 5)).astype(np.float32)
    frames = []
    for i in range(20):
        T = 9 + (7 * i) % 11
        f = scene[:T] + 0.02 * rng.standard_normal((T, 5)).astype(np.float32)
        frames.append((np.clip(f[:, :4], 0, 1), np.clip(f[:, 4:], 0, 1)))
    kw = dict(img_size=(1080, 1920), n_frames=n_frames, inactive_patience=2)
    a = KineSequenceTracker(det, use_graphs=True, token_bucket=8, track_bucket=4, **kw)
    b = KineSequenceTracker(det, use_graphs=False, **kw)
    with torch.no_grad():
        sc = b._detect_eager(frames[0], None, None)[:, 0]
    thr = float(torch.quantile(sc, 0.6))
    for t in (a, b):
        t.det_thresh = thr
    started = 0
    for i, f in enumerate(frames):
        for t in (a, b):                       # near-binary scores with random weights: drive terminations / returns by hand
            t.trk_thresh = float("inf") if i % 5 == 3 else thr - 1.0
            t.reid_thresh = float("-inf") if i % 5 == 4 else float("inf")
        ra, rb = a.step(f), b.step(f)
        assert sorted(ra) == sorted(rb) and a.inactive_tracks == b.inactive_tracks
        for tid in ra:
            assert np.abs(ra[tid]["bbox"] - rb[tid]["bbox"]).max() < 1e-3 and abs(ra[tid]["score"] - rb[tid]["score"]) < 1e-5
        started = max(started, len(ra))
    assert started > 0 and a.num_reids == b.num_reids > 0 and a.next_id == b.next_id
    assert len(a._graphs.recorded_buckets) >= 3 and all(len(k) == 6 for k in a._graphs.recorded_buckets)


@pytest.mark.gpu
def test_kine_tracker_survives_a_frame_without_detections():
    """T = 0: the graph path pads to one fully masked token bucket (attention over no keys gives zeros); live tracks are
    still queried and updated."""
    from kinet_b200.modules import kinematic_transformer as kt
    from kinet_b200.modules.kinematic_detector import KinematicDetectorTransformer
    from kinet_b200.tracking import KineSequenceTracker
    dev = torch.device("cuda:0")
    d, nq, n_frames = 72, 12, 5
    transformer = kt.DualKinematicTransformer(d_model=d, nhead=8, num_encoder_layers=1, num_decoder_layers=1,
                                              dim_feedforward=48, dropout=0.0, return_intermediate_dec=True)
    det = _by_name(KinematicDetectorTransformer(_backbones(dev, max_det=60), transformer, num_classes=1, num_queries=nq,
                                                dim_tracklets_det=4 * n_frames, dim_tracklets_metadata=n_frames)).to(dev).eval()
    trk = KineSequenceTracker(det, img_size=(1080, 1920), n_frames=n_frames, use_graphs=True, token_bucket=8, track_bucket=4,
                              detection_obj_score_thresh=-1.0, track_obj_score_thresh=-1.0)
    rng = np.random.default_rng(0)
    full = (rng.random((9, 4)).astype(np.float32), rng.random((9, 1)).astype(np.float32))
    empty = (np.zeros((0, 4), np.float32), np.zeros((0, 1), np.float32))
    a = trk.step(full)
    b = trk.step(empty)
    c = trk.step(full)
    assert len(a) > 0 and len(b) >= len(a) and len(c) >= len(b)
    for r in list(b.values()) + list(c.values()):
        assert np.isfinite(r["bbox"]).all() and np.isfinite(r["score"])
