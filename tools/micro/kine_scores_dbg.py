import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "tests"))
import numpy as np, torch
import test_kine as T
from kinet_b200.modules import kinematic_transformer as kt
from kinet_b200.modules.kinematic_detector import KinematicDetectorTransformer, TokenFeatures
from kinet_b200.tracking import KineSequenceTracker
dev = torch.device("cuda:0")
d, nq, n_frames = 72, 12, 5
back = [lambda x: ([TokenFeatures(x[0], torch.zeros(x[0].shape[:2], dtype=torch.bool, device=x[0].device))], [x[1]])] * 2
transformer = kt.DualKinematicTransformer(d_model=d, nhead=8, num_encoder_layers=1, num_decoder_layers=1, dim_feedforward=48, dropout=0.0, return_intermediate_dec=True)
det = T._by_name(KinematicDetectorTransformer(back, transformer, num_classes=1, num_queries=nq, aux_loss=False, dim_tracklets_det=4 * n_frames, dim_tracklets_metadata=n_frames)).to(dev).eval()
with torch.no_grad():
    det.class_embed.weight.mul_(6.0); det.bbox_embed.layers[-1].weight.mul_(8.0)
g = torch.Generator().manual_seed(21)
scene = [torch.randn(1, 20, d, generator=g) for _ in range(4)]
frames = []
for i in range(24):
    Tn = 9 + (7 * i) % 11
    frames.append(tuple((s[:, :Tn] + 0.25 * torch.randn(1, Tn, d, generator=g)).to(dev) for s in scene))
for trk_off, reid_off in ((-0.02, -0.02), (0.05, -0.05), (0.1, -0.1), (0.2, -0.1)):
    b = KineSequenceTracker(det, img_size=(1080, 1920), n_frames=n_frames, detection_nms_thresh=0.9, track_nms_thresh=0.9, inactive_patience=3)
    with torch.no_grad():
        sc = b._detect_eager(frames[0], None, None)[:, 0]
    thr = float(torch.quantile(sc, 0.67)); rng = float(sc.max()) - float(sc.min())
    b.det_thresh, b.trk_thresh, b.reid_thresh = thr, thr + trk_off * rng, thr + reid_off * rng
    log = []
    for f in frames:
        r = b.step(f)
        log.append((len(r), len(b.inactive_tracks)))
    print(trk_off, reid_off, "started", b.next_id, "reids", b.num_reids, log)
