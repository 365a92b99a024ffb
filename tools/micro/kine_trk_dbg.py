import sys, os, zlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from kinet_b200.modules import kinematic_transformer as kt
from kinet_b200.modules.kinematic_detector import KinematicDetectorTransformer, TokenFeatures
from kinet_b200.tracking import KineSequenceTracker
dev = torch.device("cuda:0")
d, nq, n_frames = 72, 12, 5
back = [lambda x: ([TokenFeatures(x[0], torch.zeros(x[0].shape[:2], dtype=torch.bool, device=x[0].device))], [x[1]])] * 2
torch.manual_seed(0)
transformer = kt.DualKinematicTransformer(d_model=d, nhead=8, num_encoder_layers=1, num_decoder_layers=1,
                                          dim_feedforward=48, dropout=0.0, return_intermediate_dec=True)
det = KinematicDetectorTransformer(back, transformer, num_classes=1, num_queries=nq, aux_loss=False,
                                   dim_tracklets_det=4 * n_frames, dim_tracklets_metadata=n_frames).to(dev).eval()
g = torch.Generator().manual_seed(21)
scene = [torch.randn(1, 20, d, generator=g) for _ in range(4)]
a = KineSequenceTracker(det, img_size=(1080, 1920), n_frames=n_frames, use_graphs=True, token_bucket=8, track_bucket=4)
def state(tag, T):
    for key, (gr, bufs, out) in a._graphs._graphs.items():
        print("   ", tag, key, "block", bufs["block"][:2].tolist(), "mask", bufs["mask"].int().tolist(), "kpad", bufs["kpad"].int().tolist(),
              "pad rows", float(bufs["src_det"][:, T:].abs().max()) if bufs["src_det"].shape[1] > T else None)
for K in (0, 5):
    for T in (9, 13):
        f = tuple((s[:, :T] + 0.25 * torch.randn(1, T, d, generator=g)).to(dev) for s in scene)
        td = np.random.default_rng(K).random((K, 20)).astype(np.float32)
        tm = np.random.default_rng(K + 1).random((K, 5)).astype(np.float32)
        with torch.no_grad():
            pg = a._detect_graph(f, td, tm).clone(); state("after 1st", T)
            pg1 = a._detect_graph(f, td, tm).clone(); state("after 2nd", T)
            pe = a._detect_eager(f, torch.from_numpy(td).to(dev) if K else None, torch.from_numpy(tm).to(dev) if K else None)
            pg2 = a._detect_graph(f, td, tm).clone()
        print(K, T, float((pg - pe).abs().max()), float((pg1 - pe).abs().max()), float((pg2 - pe).abs().max()))
