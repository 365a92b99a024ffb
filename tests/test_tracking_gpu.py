"""The per-frame tracking loop (kinet_b200/tracking.py; reference src/trackformer/models/tracker.py:269-557,
src/track.py:127-161) on synthetic sequences: frame t's surviving detections are frame t+1's track queries, so the
number of decoder queries changes every frame.  ``-m gpu``.

Parity (model level): the same loop with this library's kernels and with the REFERENCE's own
CUDA kernels (oracle/_ref, compiled from /root/reference for sm_100a) as the sampling op must produce identical track
ids on every frame and boxes within 1e-3 px (fp32).  Scores within 1e-4 of a threshold would make the comparison a
coin toss between two correct implementations; the test checks that the sequence has none.
"""
import numpy as np
import pytest
import torch

from _harness import RefCudaFunction, randomise_by_name

pytestmark = pytest.mark.gpu

LEVELS = [(24, 40), (12, 20), (6, 10)]
IMG = (384, 640)
D, NQ = 128, 40


def _model(dev, seed_shift=0):
    from kinet_b200.modules import DeformableTransformer, MSDeformAttn
    from kinet_b200.tracking import DetectionHeads
    torch.manual_seed(0)
    tr = DeformableTransformer(D, 8, 2, 2, 256, 0.0, "relu", True, len(LEVELS), 4, 4)
    heads = DetectionHeads(D, 1)
    randomise_by_name(tr)
    randomise_by_name(heads)
    with torch.no_grad():
        heads.class_embed.bias.zero_()                 # scores straddle 0.5: detections come and go
        heads.class_embed.weight.mul_(3.0)
        for m in tr.modules():
            if isinstance(m, MSDeformAttn):
                m._reset_parameters()                  # compass offsets
                g = torch.Generator().manual_seed(5)
                m.sampling_offsets.weight.copy_(torch.randn(m.sampling_offsets.weight.shape, generator=g) * 0.02)
                m.attention_weights.weight.copy_(torch.randn(m.attention_weights.weight.shape, generator=g) * 0.1)
    g = torch.Generator().manual_seed(11)
    query_embed = torch.randn(NQ, 2 * D, generator=g).to(dev)
    return tr.to(dev).eval(), heads.to(dev).eval(), query_embed


def _run(tracker, n_frames, seed=0):
    from kinet_b200.tracking import synthetic_sequence
    frames, counts, margins = [], [], []
    for srcs, masks, pos in synthetic_sequence(LEVELS, D, n_frames, torch.device("cuda:0"), seed=seed, drift=0.25):
        counts.append(len(tracker.track_ids) + tracker.num_object_queries)
        frames.append(tracker.step(srcs, masks, pos))
    return frames, counts


def _use_reference_kernels(tr, on):
    """Swap the sampling op of every MSDeformAttn for the reference's kernels (PyTorch prologue, fp32)."""
    import kinet_b200.modules.ms_deform_attn as mod
    from kinet_b200.modules import MSDeformAttn
    for m in tr.modules():
        if isinstance(m, MSDeformAttn):
            m.fused_prologue = not on
    if on:
        _use_reference_kernels.saved = mod.MSDeformAttnFunction
        mod.MSDeformAttnFunction = RefCudaFunction
    else:
        mod.MSDeformAttnFunction = _use_reference_kernels.saved


def _compare(fa, fb, tol_px):
    assert len(fa) == len(fb)
    for t, (a, b) in enumerate(zip(fa, fb)):
        assert sorted(a) == sorted(b), ("track ids differ on frame", t, sorted(a), sorted(b))
        for k in a:
            assert np.abs(a[k]["bbox"] - b[k]["bbox"]).max() <= tol_px, (t, k)
            assert abs(a[k]["score"] - b[k]["score"]) <= 1e-5


def test_tracking_loop_matches_reference_kernels_24_frames():
    from kinet_b200.tracking import SequenceTracker
    dev = torch.device("cuda:0")
    tr, heads, qe = _model(dev)
    kw = dict(img_size=IMG, detection_obj_score_thresh=0.5, track_obj_score_thresh=0.5)
    ours = SequenceTracker(tr, heads, qe, **kw)
    fa, counts = _run(ours, 24)
    # the sequence exercises the loop: tracks exist, are born and die, the query count keeps changing
    assert max(counts) > NQ + 3 and len(set(counts)) >= 5, counts
    assert len(ours.results) > max(c - NQ for c in counts), "no track ever ended"
    _use_reference_kernels(tr, True)
    try:
        ref = SequenceTracker(tr, heads, qe, **kw)
        fb, counts_b = _run(ref, 24)
    finally:
        _use_reference_kernels(tr, False)
    assert counts == counts_b
    _compare(fa, fb, 1e-3)


def test_tracking_loop_cuda_graphs_match_eager():
    """FrameGraphs (bucketed, mask-padded decoder graphs) against the eager loop: same ids, same boxes (fp32); and a
    bf16-autocast graph run tracks the same objects on most frames (different arithmetic: statistics only)."""
    from kinet_b200.tracking import SequenceTracker
    dev = torch.device("cuda:0")
    tr, heads, qe = _model(dev)
    kw = dict(img_size=IMG, detection_obj_score_thresh=0.5, track_obj_score_thresh=0.5)
    fa, counts = _run(SequenceTracker(tr, heads, qe, **kw), 20)
    graphed = SequenceTracker(tr, heads, qe, use_graphs=True, bucket=8, **kw)
    fb, _ = _run(graphed, 20)
    _compare(fa, fb, 1e-3)
    assert len(graphed._graphs.recorded_buckets) >= 2          # the query count crossed bucket boundaries
    # a second sequence through the same tracker object reuses the recorded graphs
    graphed.reset()
    fc, _ = _run(graphed, 20)
    _compare(fa, fc, 1e-3)
    # the pipelined generator (next frame's encoder enqueued before the read-back) gives the same results
    from kinet_b200.tracking import synthetic_sequence
    graphed.reset()
    fp = list(graphed.run(synthetic_sequence(LEVELS, D, 20, dev, seed=0, drift=0.25)))
    _compare(fa, fp, 1e-3)
    lowp = SequenceTracker(tr, heads, qe, use_graphs=True, bucket=8, autocast_dtype=torch.bfloat16, **kw)
    fd, _ = _run(lowp, 20)
    same = sum(sorted(a) == sorted(b) for a, b in zip(fa, fd))
    assert len(fd) == 20 and all(len(f) > 0 for f in fd[1:])
    print("bf16 graphs: identical id sets on %d of 20 frames" % same)


def test_tracker_termination_and_nms_rules_on_host():
    """The host-side rules in isolation (no GPU work needed beyond tensors): NMS keeps existing tracks over new
    detections, a track below the threshold survives steps_termination - 1 frames with its old box."""
    from kinet_b200.tracking import _nms
    boxes = np.array([[0, 0, 10, 10], [0.5, 0.5, 10.5, 10.5], [50, 50, 60, 60.]])
    assert _nms(boxes, np.array([0.6, np.inf, 0.9]), 0.5).tolist() == [1, 2]
    assert _nms(boxes, np.array([0.9, 0.8, 0.7]), 0.95).tolist() == [0, 1, 2]


def test_tracking_loop_inactive_tracks_return_and_graphs_match_eager():
    """inactive_patience >= 0 (tracker.py:273-276, :290-304, :367-394): terminated tracks stay track queries (last box +
    embedding) behind the live ones for that many frames and return under their id when their query scores above
    reid_score_thresh.  Eager loop, bucketed graphs and the pipelined generator agree frame by frame; with the default
    (-1) nothing is ever inactive."""
    from kinet_b200.tracking import SequenceTracker, synthetic_sequence
    dev = torch.device("cuda:0")
    tr, heads, qe = _model(dev)
    kw = dict(img_size=IMG, detection_obj_score_thresh=0.5, track_obj_score_thresh=0.5, inactive_patience=3,
              reid_score_thresh=0.45)

    def run(tracker, pipelined=False):
        seq = synthetic_sequence(LEVELS, D, 24, dev, seed=0, drift=0.25)
        frames, inactive = [], []
        if pipelined:
            for f in tracker.run(seq):
                frames.append(f)
                inactive.append(list(tracker.inactive_ids))
        else:
            for srcs, masks, pos in seq:
                frames.append(tracker.step(srcs, masks, pos))
                inactive.append(list(tracker.inactive_ids))
        return frames, inactive

    eager = SequenceTracker(tr, heads, qe, **kw)
    fa, ia = run(eager)
    assert any(ia), "no track ever went inactive"
    assert all(c <= 3 + 1 for c in eager.inactive_count)
    assert eager.num_reids > 0, "no inactive track ever returned"
    # a returned track has a gap in its results: present, absent while inactive, present again under the same id
    gaps = [t for t, r in eager.results.items() if len(r) >= 2 and max(r) - min(r) + 1 > len(r)]
    assert gaps
    # inactive tracks are not reported while inactive
    for f, ids in zip(fa, ia):
        assert not (set(f) & set(ids))
    graphed = SequenceTracker(tr, heads, qe, use_graphs=True, bucket=8, **kw)
    fb, ib = run(graphed)
    _compare(fa, fb, 1e-3)
    assert ia == ib and graphed.num_reids == eager.num_reids
    graphed.reset()
    fc, ic = run(graphed, pipelined=True)
    _compare(fa, fc, 1e-3)
    assert ia == ic
    plain = SequenceTracker(tr, heads, qe, **dict(kw, inactive_patience=-1))
    fd, idd = run(plain)
    assert not any(idd) and plain.num_reids == 0


def test_tracking_loop_public_detections_gate_births():
    """public_detections (tracker.py:126-167, :446-448): without a backing public detection no track is born; fed the plain
    run's own boxes as public detections, the gated loop (eager and pipelined graphs) reproduces the plain run."""
    from kinet_b200.tracking import SequenceTracker, synthetic_sequence
    dev = torch.device("cuda:0")
    tr, heads, qe = _model(dev)
    kw = dict(img_size=IMG, detection_obj_score_thresh=0.5, track_obj_score_thresh=0.5)
    plain, _ = _run(SequenceTracker(tr, heads, qe, **kw), 12)
    gated = SequenceTracker(tr, heads, qe, public_detections="min_iou_0_5", **kw)
    seq = list(synthetic_sequence(LEVELS, D, 12, dev, seed=0, drift=0.25))
    none = [gated.step(*f) for f in seq]
    assert all(len(f) == 0 for f in none) and gated.next_id == 0
    pub = [np.stack([r["bbox"] for r in f.values()]) if f else None for f in plain]
    gated.reset()
    fa = [gated.step(*f, public_dets=p) for f, p in zip(seq, pub)]
    _compare(plain, fa, 1e-3)
    graphed = SequenceTracker(tr, heads, qe, public_detections="min_iou_0_5", use_graphs=True, bucket=8, **kw)
    fb = list(graphed.run([f + (p,) for f, p in zip(seq, pub)]))
    _compare(plain, fb, 1e-3)


@pytest.mark.parametrize("greedy", [True, False])
def test_tracking_loop_reid_of_new_detections_with_inactive_tracks(greedy):
    """reid() (tracker.py:170-267): with the query-based return switched off, inactive tracks come back only by claiming a
    new detection -- greedily by centre distance, or by Hungarian assignment on the embedding distance below
    reid_sim_threshold.  A claimed detection starts no track of its own; eager and graphs agree."""
    from kinet_b200.tracking import SequenceTracker, synthetic_sequence
    dev = torch.device("cuda:0")
    tr, heads, qe = _model(dev)
    kw = dict(img_size=IMG, detection_obj_score_thresh=0.5, track_obj_score_thresh=0.5, inactive_patience=4,
              reid_score_thresh=float("inf"), reid_greedy_matching=greedy, reid_sim_threshold=0.0 if greedy else 1e9)
    seq = list(synthetic_sequence(LEVELS, D, 24, dev, seed=0, drift=0.25))
    eager = SequenceTracker(tr, heads, qe, **kw)
    fa = [eager.step(*f) for f in seq]
    assert eager.num_reids > 0
    plain = SequenceTracker(tr, heads, qe, **dict(kw, inactive_patience=-1))
    fp = [plain.step(*f) for f in seq]
    assert plain.num_reids == 0 and eager.next_id < plain.next_id          # returned tracks took detections that were births
    graphed = SequenceTracker(tr, heads, qe, use_graphs=True, bucket=8, **kw)
    fb = list(graphed.run(seq))
    _compare(fa, fb, 1e-3)
    assert graphed.num_reids == eager.num_reids
