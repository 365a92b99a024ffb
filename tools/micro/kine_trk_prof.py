"""cProfile of KineSequenceTracker._update on recorded detector outputs of the kine_bench tracker loop."""
import sys, os, time, cProfile, pstats
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import kine_bench
from kinet_b200 import tracking
rec = []
thr = [0, 0]
orig = tracking.KineSequenceTracker._update
def spy(self, packed, K):
    rec.append((packed.copy(), K, [a.copy() for a in (self.ids, self.boxes, self.scores, self.misses, self.trail, self.trail_sc)], self.next_id))
    thr[:] = [self.det_thresh, self.trk_thresh]
    return orig(self, packed, K)
tracking.KineSequenceTracker._update = spy
res = kine_bench.measure_tracker(torch.device("cuda:0"), frames=100)
tracking.KineSequenceTracker._update = orig
print(res)
rec = rec[-100:]
trk = tracking.KineSequenceTracker(None, (1080, 1920), n_frames=5)
trk.det_thresh, trk.trk_thresh = thr
def replay():
    for packed, K, st, nid in rec:
        trk.ids, trk.boxes, trk.scores, trk.misses, trk.trail, trk.trail_sc = st
        trk.next_id = nid
        trk._update(packed, K)
# thresholds as measured
import json
replay()
t = time.perf_counter(); replay(); print("update us/frame ", (time.perf_counter() - t) / len(rec) * 1e6)
np.save("gpurun_out/kine_rec_packed.npy", np.concatenate([r[0] for r in rec[:20]]))
pr = cProfile.Profile(); pr.enable(); replay(); pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(18)
