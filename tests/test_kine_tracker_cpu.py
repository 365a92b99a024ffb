"""Host logic of kinet_b200.tracking.KineSequenceTracker (the loop of the reference's TrackerKinematic.step,
src/trackformer/models/tracker.py:626-880) on hand-made detector outputs: births, trails, misses and terminations,
the two suppression passes (tracks among themselves; old tracks always win against births).  No GPU, no detector."""
import numpy as np

from kinet_b200.tracking import FrameTracks, KineSequenceTracker, _greedy_keep, _iou_matrix, _nms

W, H = 1000.0, 500.0


def _tracker(**kw):
    args = dict(img_size=(H, W), n_frames=3, detection_obj_score_thresh=0.5, track_obj_score_thresh=0.4,
                detection_nms_thresh=0.6, track_nms_thresh=0.6, steps_termination=1)
    args.update(kw)
    return KineSequenceTracker(None, **args)


def _row(score, cx, cy, w=0.1, h=0.2, label=0):
    return [score, label, cx, cy, w, h]


def _packed(rows):
    return np.asarray(rows, np.float32).reshape(-1, 6)


def test_births_pixel_boxes_and_pseudo_trails():
    t = _tracker()
    out = t._update(_packed([_row(0.9, 0.5, 0.5), _row(0.3, 0.2, 0.2), _row(0.7, 0.1, 0.8), _row(0.95, 0.7, 0.3, label=1)]), 0)
    assert isinstance(out, FrameTracks) and sorted(out) == [0, 1] and t.next_id == 2      # below threshold / wrong class: no track
    np.testing.assert_allclose(out[0]["bbox"], [450, 200, 550, 300], atol=1e-4)       # cxcywh relative -> xyxy pixels
    np.testing.assert_allclose(out[1]["bbox"], [50, 350, 150, 450], atol=1e-4)
    assert abs(out[1]["score"] - 0.7) < 1e-6
    # generate_pseudo_tracklets: the detection repeated n_frames times
    np.testing.assert_allclose(t.trail[0], np.tile([0.5, 0.5, 0.1, 0.2], (3, 1)), atol=1e-6)
    np.testing.assert_allclose(t.trail_sc[1], [0.7, 0.7, 0.7], atol=1e-6)
    det, meta = t._trails()
    assert det.shape == (2, 12) and meta.shape == (2, 3)


def test_track_survives_shifts_its_trail_and_a_miss_ends_it():
    t = _tracker()
    t._update(_packed([_row(0.9, 0.5, 0.5), _row(0.7, 0.1, 0.8)]), 0)
    # track 0 re-detected a little to the right, track 1 falls below the track threshold; one object query fires far away
    out = t._update(_packed([_row(0.8, 0.52, 0.5), _row(0.2, 0.1, 0.8), _row(0.6, 0.85, 0.2)]), 2)
    assert sorted(out) == [0, 2] and t.next_id == 3
    np.testing.assert_allclose(out[0]["bbox"], [470, 200, 570, 300], atol=1e-4)
    np.testing.assert_allclose(t.trail[0], [[0.5, 0.5, 0.1, 0.2], [0.5, 0.5, 0.1, 0.2], [0.52, 0.5, 0.1, 0.2]], atol=1e-6)
    np.testing.assert_allclose(t.trail_sc[0], [0.9, 0.9, 0.8], atol=1e-6)
    assert t.results[0].keys() == {0, 1} and t.results[1].keys() == {0} and t.results[2].keys() == {1}


def test_a_track_with_a_foreign_label_counts_as_missed_and_patience_keeps_the_old_box():
    t = _tracker(steps_termination=2)
    t._update(_packed([_row(0.9, 0.5, 0.5)]), 0)
    out = t._update(_packed([_row(0.9, 0.6, 0.6, label=1), _row(0.0, 0, 0)]), 1)          # best class is not class 0
    assert sorted(out) == [0] and t.misses.tolist() == [1]
    np.testing.assert_allclose(out[0]["bbox"], [450, 200, 550, 300], atol=1e-4)           # position of the last hit
    np.testing.assert_allclose(t.trail[0], np.tile([0.5, 0.5, 0.1, 0.2], (3, 1)), atol=1e-6)
    out = t._update(_packed([_row(0.9, 0.6, 0.6), _row(0.0, 0, 0)]), 1)                    # found again: misses reset
    assert t.misses.tolist() == [0]
    np.testing.assert_allclose(out[0]["bbox"], [550, 250, 650, 350], atol=1e-4)
    t._update(_packed([_row(0.1, 0.6, 0.6), _row(0.0, 0, 0)]), 1)
    out = t._update(_packed([_row(0.1, 0.6, 0.6), _row(0.0, 0, 0)]), 1)                    # second miss in a row: gone
    assert len(out) == 0 and len(t.ids) == 0


def test_existing_tracks_win_against_overlapping_births_and_better_births_win_among_themselves():
    t = _tracker()
    t._update(_packed([_row(0.55, 0.5, 0.5)]), 0)
    out = t._update(_packed([_row(0.45, 0.5, 0.5),            # the track, low score but above the track threshold
                             _row(0.99, 0.505, 0.5),          # a birth on top of it: loses although its score is higher
                             _row(0.6, 0.2, 0.2), _row(0.8, 0.202, 0.2),     # two births on top of each other: 0.8 wins
                             _row(0.7, 0.8, 0.8)]), 1)
    assert sorted(out) == [0, 1, 2] and t.next_id == 3
    by_score = {round(v["score"], 2) for v in out.values()}
    assert by_score == {0.45, 0.8, 0.7}


def test_track_nms_drops_the_weaker_of_two_tracks_that_converged():
    t = _tracker()
    t._update(_packed([_row(0.9, 0.3, 0.5), _row(0.8, 0.7, 0.5)]), 0)
    out = t._update(_packed([_row(0.6, 0.5, 0.5), _row(0.7, 0.501, 0.5), _row(0.0, 0, 0)]), 2)
    assert sorted(out) == [1]
    t = _tracker(track_nms_thresh=0, detection_nms_thresh=0)                                # both passes off
    t._update(_packed([_row(0.9, 0.3, 0.5), _row(0.8, 0.7, 0.5)]), 0)
    out = t._update(_packed([_row(0.6, 0.5, 0.5), _row(0.7, 0.501, 0.5), _row(0.9, 0.5, 0.5)]), 2)
    assert sorted(out) == [0, 1, 2]


def test_boxes_are_clamped_to_the_image():
    t = _tracker()
    out = t._update(_packed([_row(0.9, 0.02, 0.99, w=0.2, h=0.2), [0.9, 0, 1.4, -0.2, 0.1, 0.1]]), 0)
    np.testing.assert_allclose(out[0]["bbox"], [0, 445, 120, 500], atol=1e-3)
    np.testing.assert_allclose(out[1]["bbox"], [950, 0, 1000, 25], atol=1e-3)             # relative box clipped to [0, 1] first
    np.testing.assert_allclose(t.trail[1][0], [1.0, 0.0, 0.1, 0.1], atol=1e-6)


def test_state_arrays_are_replaced_not_edited():
    """FrameTracks of an earlier frame keeps showing that frame (the tracker stores references, the dicts are lazy)."""
    t = _tracker()
    first = t._update(_packed([_row(0.9, 0.5, 0.5)]), 0)
    t._update(_packed([_row(0.8, 0.6, 0.6), _row(0.9, 0.1, 0.1)]), 1)
    np.testing.assert_allclose(first[0]["bbox"], [450, 200, 550, 300], atol=1e-4)
    assert len(first) == 1 and abs(first[0]["score"] - 0.9) < 1e-6


def test_iou_matrix_and_nms_semantics():
    boxes = np.array([[0, 0, 10, 10], [0, 0, 10, 5], [20, 20, 30, 30], [5, 5, 5, 5]], np.float64)
    iou = _iou_matrix(boxes)
    np.testing.assert_allclose(iou[0], [1.0, 0.5, 0.0, 0.0], atol=1e-6)
    np.testing.assert_allclose(iou, iou.T, atol=0)
    assert _nms(boxes, np.array([0.5, 0.9, 0.1, 0.2]), 0.4).tolist() == [1, 2, 3]         # IoU 0.5 > 0.4: the 0.9 box wins
    assert _nms(boxes, np.array([0.5, 0.9, 0.1, 0.2]), 0.5).tolist() == [0, 1, 2, 3]      # strictly greater suppresses
    over = np.array([[0, 1, 0], [1, 0, 1], [0, 1, 0]], bool)
    assert _greedy_keep(over, [1, 0, 2]).tolist() == [1] and _greedy_keep(over, [0, 1, 2]).tolist() == [0, 2]


def _step(t, rows):
    """One frame through the host logic as KineSequenceTracker.step runs it: prune, build the query list, update."""
    t._prune_inactive()
    K = len(t.ids) + len(t.inactive[0])
    det, meta = t._trails()
    assert det.shape == (K, 4 * t.n_frames) and meta.shape == (K, t.n_frames)
    assert len(rows) >= K
    return t._update(_packed(rows), K), det, meta


def test_inactive_track_stays_a_query_and_comes_back_under_its_id():
    t = _tracker(inactive_patience=2, reid_score_thresh=0.3)
    _step(t, [_row(0.9, 0.5, 0.5), _row(0.8, 0.2, 0.2)])                        # two births: ids 0, 1
    out, _, _ = _step(t, [_row(0.9, 0.5, 0.5), _row(0.1, 0.2, 0.2), _row(0.0, 0, 0)])      # track 1 is lost
    assert sorted(out) == [0] and t.inactive_tracks == [1] and t.inactive[5].tolist() == [1]
    # its frozen trail (last entry repeated once) is the frame's second track query
    out, det, _ = _step(t, [_row(0.9, 0.5, 0.5), _row(0.2, 0.2, 0.2), _row(0.0, 0, 0)])    # still below reid_score_thresh
    np.testing.assert_allclose(det[1].reshape(3, 4), np.tile([0.2, 0.2, 0.1, 0.2], (3, 1)), atol=1e-6)
    assert sorted(out) == [0] and t.inactive_tracks == [1] and t.inactive[5].tolist() == [2]
    out, _, _ = _step(t, [_row(0.9, 0.5, 0.5), _row(0.35, 0.25, 0.2), _row(0.0, 0, 0)])    # above it (and below the track threshold)
    assert sorted(out) == [0, 1] and t.inactive_tracks == [] and t.num_reids == 1 and t.next_id == 2
    np.testing.assert_allclose(out[1]["bbox"], [200, 50, 300, 150], atol=1e-3)
    np.testing.assert_allclose(t.trail[1], [[0.2, 0.2, 0.1, 0.2], [0.2, 0.2, 0.1, 0.2], [0.25, 0.2, 0.1, 0.2]], atol=1e-6)
    assert t.results[1].keys() == {0, 3}                                                  # absent while inactive


def test_inactive_track_is_forgotten_after_its_patience_and_default_forgets_at_once():
    t = _tracker(inactive_patience=1)
    _step(t, [_row(0.9, 0.5, 0.5)])
    _step(t, [_row(0.1, 0.5, 0.5), _row(0.0, 0, 0)])                                       # lost: inactive, count 1
    assert t.inactive_tracks == [0]
    _step(t, [_row(0.1, 0.5, 0.5), _row(0.0, 0, 0)])                                       # count 1 <= patience: still a query; count 2
    assert t.inactive_tracks == [0]
    out, det, _ = _step(t, [_row(0.9, 0.5, 0.5)])                                          # pruned before the frame: one object query only
    assert det.shape[0] == 0 and t.inactive_tracks == [] and sorted(out) == [1]           # the detection starts a NEW track
    d = _tracker()                                                                         # inactive_patience = -1 (cfgs/track.yaml)
    _step(d, [_row(0.9, 0.5, 0.5)])
    _step(d, [_row(0.1, 0.5, 0.5), _row(0.0, 0, 0)])
    assert d.inactive_tracks == [] and len(d.ids) == 0


def test_a_degenerate_inactive_box_is_dropped_and_reactivated_tracks_take_part_in_nms():
    t = _tracker(inactive_patience=5, reid_score_thresh=0.3)
    _step(t, [_row(0.9, 0.5, 0.5, w=0.0), _row(0.9, 0.2, 0.2), _row(0.8, 0.7, 0.7)])       # id 0 has no area
    _step(t, [_row(0.1, 0.5, 0.5), _row(0.1, 0.2, 0.2), _row(0.9, 0.7, 0.7), _row(0.0, 0, 0)])
    assert t.inactive_tracks == [0, 1]
    t._prune_inactive()
    assert t.inactive_tracks == [1]                                                        # has_positive_area (:631)
    # track 1 comes back on top of track 2 with the lower score: NMS among the tracks removes it again
    out, _, _ = _step(t, [_row(0.9, 0.7, 0.7), _row(0.5, 0.701, 0.7), _row(0.0, 0, 0)])
    assert sorted(out) == [2] and t.num_reids == 1 and t.inactive_tracks == []


def test_public_detections_mask_and_its_use_for_births():
    """tracker.py:126-167: per public detection the nearest (centre distance below the detection's area) / best
    overlapping (IoU >= 0.5) unclaimed detection; births only where backed (:905-910)."""
    from kinet_b200.tracking import public_detections_mask as pdm
    new = np.array([[0, 0, 10, 10], [20, 20, 30, 30], [0, 0, 9, 9.]])
    pub = np.array([[1, 1, 11, 11.], [100, 100, 110, 110]])
    assert pdm(new, pub, False).tolist() == [True, True, True]
    assert pdm(new, pub, "center_distance").tolist() == [True, False, False]      # one public box claims ONE detection
    assert pdm(new, pub, "min_iou_0_5").tolist() == [True, False, False]
    assert pdm(new, np.array([[1, 1, 11, 11.], [0, 0, 9, 9]]), "min_iou_0_5").tolist() == [True, False, True]
    assert pdm(new, None, "min_iou_0_5").tolist() == [False, False, False] and pdm(new[:0], pub, "center_distance").tolist() == []
    try:
        pdm(new, pub, "other")
        assert False
    except NotImplementedError:
        pass
    t = _tracker(public_detections="min_iou_0_5")
    rows = [_row(0.9, 0.5, 0.5), _row(0.8, 0.2, 0.2)]                              # pixel boxes [450,200,550,300], [150,0,250,100]
    out = t._update(_packed(rows), 0, np.array([[445, 195, 545, 295.]]))
    assert sorted(out) == [0] and abs(out[0]["score"] - 0.9) < 1e-6
    out = t._update(_packed([_row(0.9, 0.5, 0.5)] + rows), 1, None)              # no public detections: no births, the track stays
    assert sorted(out) == [0]
