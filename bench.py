#!/usr/bin/env python
"""bench.py -- MSDeformAttn forward+backward throughput on the Deformable-DETR 6+6-layer shape.

Contract (see the task brief): `python bench.py --gpus N --steps K --warmup W [--impl reference]`
prints ONE JSON line on rank 0.

Workload ("detr_6enc_6dec_bf16", BASELINE.json configs[1]): one step = the MSDeformAttn calls of one
Deformable-DETR forward+backward pass on an 800x1333 pyramid -- 6 encoder-layer calls (N=2, Lq=S=22223,
M=8, D=32, L=4, P=4) and 6 decoder-layer calls (Lq=300), each forward AND backward, value in bf16,
sampling locations / attention weights in fp32 (what autocast produces), every layer with its own
input tensors (~0.7 GB per step > 126 MB L2, so no call finds its inputs in L2).

metric: msda_fwd_bwd_GBps = algorithmic bytes of the step / device time.
  value    inputs resident in HBM, calls made straight through the C ABI (libmsda_b200.so)
  e2e      same step through the public Python API (MSDeformAttnFunction) with HOST pinned inputs:
           H2D of every input, forward, backward, D2H of output and gradients inside the timed region
  roofline dominant kernel = the encoder-form backward (the dense-window tcgen05 kernel + the per-row kernel
           for the coarser levels: two launches the library brackets as one with CUDA events,
           msda_b200_trace_next); DRAM traffic etc. come from profiles/roofline_inputs.json, written from
           an ncu capture together with a hash of kinet_b200/csrc -- a capture of other kernels is flagged stale
  sweep    BASELINE.json configs[4] (MOT20 crowd, full-res pyramid, 300 + 500 queries): per-call times
  cpu_baseline / --impl reference: the reference's CPU-runnable path (per-level grid_sample + autograd,
           restated in oracle/msda_oracle.py because /root/reference is absent on the GPU box) on the SAME
           step (6 encoder-form + 6 decoder-form calls, full Lq, fp32), all host threads.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from kinet_b200.workloads import algorithmic_bytes, make_inputs, resolve  # noqa: E402

WORKLOAD = "detr_6enc_6dec_bf16"
ENC_TAG, DEC_TAG, LAYERS = "C2_enc", "C2_dec", 6
METRIC, UNIT = "msda_fwd_bwd_GBps", "GB/s"
ROOFLINE_INPUTS = os.path.join(ROOT, "profiles", "roofline_inputs.json")


def csrc_hash():
    """sha256 over the operator's CUDA sources (csrc/msda_*: the kernels the ncu capture is of; the callers' glue
    kernels in layer_ops.cu / small_mha.cu are not part of that launch), the key profiles/roofline_inputs.json was
    recorded under."""
    import hashlib
    h = hashlib.sha256()
    d = os.path.join(ROOT, "kinet_b200", "csrc")
    for name in sorted(os.listdir(d)):
        if name.startswith("msda_") and name.endswith((".cu", ".cuh")):
            with open(os.path.join(d, name), "rb") as f:
                h.update(name.encode() + b"\0" + f.read())
    return h.hexdigest()[:16]


def roofline_inputs():
    """DRAM traffic / reduction sectors of the dominant kernel from the committed ncu capture (bench.py never runs
    under a profiler).  stale = the capture was taken on other kernel sources than the ones being measured."""
    try:
        with open(ROOFLINE_INPUTS) as f:
            r = json.load(f)
    except Exception:
        return {"traffic": None, "stale": None, "source": None}
    r["stale"] = r.get("csrc_hash") != csrc_hash()
    return r


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def step_bytes(e_value, e_aux):
    """Algorithmic bytes of one step and of one encoder-form backward call."""
    enc, dec = resolve(ENC_TAG), resolve(DEC_TAG)
    ef, eb = algorithmic_bytes(enc, e_value, e_aux)
    df, db = algorithmic_bytes(dec, e_value, e_aux)
    return LAYERS * (ef + eb + df + db), eb


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------
class Call:
    """One MSDeformAttn call site (one layer): device tensors + preallocated outputs, launched
    straight through the C ABI (no allocation, no host sync in the timed region)."""

    def __init__(self, tag, dev, seed):
        from kinet_b200 import _lib
        self.lib = _lib.load()
        self._lib = _lib
        s = resolve(tag)
        self.shape = s
        self.value, self.shapes, self.loc, self.attn = make_inputs(
            tag, dev, torch.bfloat16, torch.float32, "structured", seed)
        g = torch.Generator(device=dev).manual_seed(seed + 1000)
        self.grad_out = torch.randn(s["N"], s["Lq"], s["M"] * s["D"], generator=g, device=dev).to(torch.bfloat16)
        self.out = torch.empty_like(self.grad_out)
        self.gv = torch.empty_like(self.value)
        self.gl = torch.empty_like(self.loc)
        self.ga = torch.empty_like(self.attn)
        self.dims = _lib.MsdaDims(s["N"], s["S"], s["M"], s["D"], s["L"], s["Lq"], s["P"])
        self.vcode, self.acode = _lib.dtype_code(self.value.dtype), _lib.dtype_code(self.loc.dtype)
        nbytes = int(self.lib.msda_b200_backward_workspace_bytes(self.vcode, ctypes.byref(self.dims)))
        self.ws_bytes = nbytes
        self.ws = torch.empty(max(nbytes // 4, 1), dtype=torch.float32, device=dev)
        self.launches = 0

    @staticmethod
    def _p(t):
        return ctypes.c_void_p(t.data_ptr())

    def forward(self, stream):
        rc = self.lib.msda_b200_forward(self._p(self.value), self.vcode, self._p(self.shapes), None,
                                        self._p(self.loc), self._p(self.attn), self.acode,
                                        self._p(self.out), ctypes.byref(self.dims), ctypes.c_void_p(stream))
        self._lib.check(rc, "msda_b200_forward")
        self.launches += self.lib.msda_b200_last_launch_count()

    def backward(self, stream):
        rc = self.lib.msda_b200_backward(self._p(self.value), self.vcode, self._p(self.shapes), None,
                                         self._p(self.loc), self._p(self.attn), self.acode,
                                         self._p(self.grad_out), self._p(self.gv), self._p(self.gl),
                                         self._p(self.ga), self._p(self.ws), ctypes.c_size_t(self.ws_bytes),
                                         ctypes.byref(self.dims), ctypes.c_void_p(stream))
        self._lib.check(rc, "msda_b200_backward")
        self.launches += self.lib.msda_b200_last_launch_count()


def run_ours(args, rank, world, dev):
    from kinet_b200 import _lib
    lib = _lib.load()
    stream = torch.cuda.current_stream().cuda_stream
    enc = [Call(ENC_TAG, dev, 100 * rank + i) for i in range(LAYERS)]
    dec = [Call(DEC_TAG, dev, 100 * rank + 50 + i) for i in range(LAYERS)]
    calls = enc + dec
    total_bytes, enc_bwd_bytes = step_bytes(2, 4)

    # events bracketing the dominant kernel (encoder-form backward), one pair per launch
    n_tr = args.steps * LAYERS
    tr = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_tr)]

    def step(trace_base=None):
        # forward through the 6 encoder then 6 decoder layers, backward in reverse order
        for c in calls:
            c.forward(stream)
        for i, c in enumerate(reversed(calls)):
            if trace_base is not None and c.shape["encoder"]:
                a, b = tr[trace_base + (i - LAYERS)]
                # torch events are created lazily: record once so the handle exists, then hand the
                # raw cudaEvent_t to the library, which re-records it around the kernel
                a.record(); b.record()
                lib.msda_b200_trace_next(ctypes.c_void_p(a.cuda_event), ctypes.c_void_p(b.cuda_event))
            c.backward(stream)

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    for c in calls:
        c.launches = 0
    barrier()
    sampler = ClockSampler(dev.index or 0)
    if rank == 0:
        sampler.start()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for k in range(args.steps):
        step(trace_base=k * LAYERS)
    t1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = t0.elapsed_time(t1)
    kern_ms = [a.elapsed_time(b) for a, b in tr]
    launches = sum(c.launches for c in calls)

    # ---- e2e: public Python API, host pinned inputs, H2D + fwd + bwd + D2H per step ------------
    from kinet_b200.functions import MSDeformAttnFunction
    host = []
    for c in calls:
        host.append(tuple(t.cpu().pin_memory() for t in (c.value, c.loc, c.attn, c.grad_out)))
    h2d = sum(t.numel() * t.element_size() for tup in host for t in tup)
    d2h = sum(t.numel() * t.element_size() for c in calls for t in (c.out, c.gv, c.gl, c.ga))
    sink = [tuple(torch.empty_like(t, device="cpu").pin_memory() for t in (c.out, c.gv, c.gl, c.ga)) for c in calls]

    # device-side landing buffers for the inputs (allocated once; refilled from the host every step)
    # two sets: step k+1's inputs stream in while step k still computes and streams its results out
    lands = [[tuple(torch.empty_like(t, device=dev) for t in tup) for tup in host] for _ in range(2)]
    done = [None, None]            # event: the kernels of the step that last used this set have finished
    step_no = [0]
    cur = torch.cuda.current_stream()
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()   # PCIe is full duplex: H2D and D2H overlap

    # Results stay referenced until their device -> host copies have run (two steps later) and the compute stream has
    # waited for them; then their blocks go back to the allocator.  (Tensor.record_stream does the same job, but every
    # later allocation then polls the pending events: 80-120 us per torch.empty on this box, 5 of 9 ms of host time per step.)
    hold, d2h_done = [None, None], [None, None]

    def e2e_step():
        # copy engine (s_in): this step's inputs, layer by layer; compute follows layer by layer
        if d2h_done[step_no[0] & 1] is not None:
            cur.wait_event(d2h_done[step_no[0] & 1])     # step k-2's results are on the host
        hold[step_no[0] & 1] = None
        keep = []
        land = lands[step_no[0] & 1]
        if done[step_no[0] & 1] is not None:
            s_in.wait_event(done[step_no[0] & 1])   # kernels of step k-2 are done with this landing set
        ready = []
        with torch.cuda.stream(s_in):
            for (hv, hl, ha, hg), (dv, dl, da, dg) in zip(host, land):
                dv.copy_(hv, non_blocking=True); dl.copy_(hl, non_blocking=True)
                da.copy_(ha, non_blocking=True); dg.copy_(hg, non_blocking=True)
                ev = torch.cuda.Event(); ev.record(s_in); ready.append(ev)
        outs = []
        for c, (dv, dl, da, dg), ev in zip(calls, land, ready):
            cur.wait_event(ev)
            v = dv.detach().requires_grad_(True)
            l = dl.detach().requires_grad_(True)
            a = da.detach().requires_grad_(True)
            o = MSDeformAttnFunction.apply(v, c.shapes, l, a, 64)
            outs.append((o, v, l, a, dg))
        for (o, v, l, a, g), dst in zip(reversed(outs), reversed(sink)):
            o.backward(g)
            ev = torch.cuda.Event(); ev.record(cur)
            s_out.wait_event(ev)
            srcs = (o.detach(), v.grad, l.grad, a.grad)
            with torch.cuda.stream(s_out):
                for src, d in zip(srcs, dst):
                    d.copy_(src, non_blocking=True)
            keep.append((srcs, o, v, l, a))
        ev = torch.cuda.Event(); ev.record(cur)
        done[step_no[0] & 1] = ev
        ev = torch.cuda.Event(); ev.record(s_out)
        d2h_done[step_no[0] & 1] = ev
        hold[step_no[0] & 1] = keep
        step_no[0] += 1

    # exactly K steps like the device-resident leg (the software pipeline's drain -- the last step's results have no
    # next step's inputs to overlap with -- is inside the timed region and is amortised over them)
    e2e_steps = max(1, min(args.steps, args.e2e_steps) if args.e2e_steps > 0 else args.steps)
    # warm-up: the caching allocator reaches its steady state (outputs handed to the D2H stream with record_stream come
    # back late, so the first steps still grow the pool with cudaMalloc -- a device-wide stall inside a timed step)
    for _ in range(max(3, min(args.warmup, 4))):
        e2e_step()
    cur.wait_stream(s_in)
    cur.wait_stream(s_out)
    torch.cuda.synchronize()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s_in.wait_event(e0)                # no copy of the timed steps starts before the start mark
    prof = None
    if os.environ.get("BENCH_E2E_PROFILE"):          # development aid: where the host's enqueue time goes
        import cProfile
        prof = cProfile.Profile()
        prof.enable()
    host_t = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    if prof is not None:
        import pstats
        prof.disable()
        pstats.Stats(prof, stream=sys.stderr).sort_stats("tottime").print_stats(25)
    host_ms = (time.perf_counter() - host_t) * 1e3 / e2e_steps     # what the host needs to ENQUEUE a step
    cur.wait_stream(s_in)
    cur.wait_stream(s_out)             # the timed region ends when the last step's results are on the host
    e1.record()
    barrier()
    e2e_ms = e0.elapsed_time(e1) / e2e_steps
    ceiling_ms = copy_ceiling(host, sink, dev, world)
    return dict(ms=ms, kern_ms=kern_ms, launches=launches, clocks=clocks, total_bytes=total_bytes,
                enc_bwd_bytes=enc_bwd_bytes, e2e_ms=e2e_ms, h2d=h2d, d2h=d2h, e2e_steps=e2e_steps,
                ceiling_ms=ceiling_ms, e2e_host_ms=host_ms)


# ---------------------------------------------------------------------------------------------
# CPU arm: the reference's CPU-runnable path (grid_sample per level + autograd), bounded sample
# ---------------------------------------------------------------------------------------------
def cpu_sample_inputs(tag, lq_cap=None):
    s = resolve(tag)
    value, shapes, loc, attn = make_inputs(tag, torch.device("cpu"), torch.float32, torch.float32, "structured", 7)
    if lq_cap is not None and s["Lq"] > lq_cap:
        loc, attn = loc[:, :lq_cap].contiguous(), attn[:, :lq_cap].contiguous()
        s = dict(s, Lq=lq_cap)
    go = torch.randn(s["N"], s["Lq"], s["M"] * s["D"])
    return s, value, shapes, loc, attn, go


def cpu_step(cases):
    """forward + backward of every case through the oracle's torch restatement of the reference."""
    from oracle import torch_port_forward
    for s, value, shapes, loc, attn, go in cases:
        v = value.clone().requires_grad_(True)
        l = loc.clone().requires_grad_(True)
        a = attn.clone().requires_grad_(True)
        out = torch_port_forward(v, shapes, l, a)
        out.backward(go)


def cpu_cases():
    # the SAME step as the GPU arm: 6 encoder-form + 6 decoder-form calls, full Lq, forward + backward, fp32 (the
    # reference has no bf16); the layers share one set of inputs (the CPU has no cache that could hold them)
    enc, dec = cpu_sample_inputs(ENC_TAG), cpu_sample_inputs(DEC_TAG)
    cases = [enc] * LAYERS + [dec] * LAYERS
    nbytes = 0
    for s, *_ in cases:
        f, b = algorithmic_bytes(s, 4, 4)
        nbytes += f + b
    return cases, nbytes, ("the full step: %d encoder-form calls (N=2, Lq=S=22223) + %d decoder-form calls (Lq=300), "
                           "fwd+bwd, fp32, 800x1333 pyramid" % (LAYERS, LAYERS))


def run_cpu(steps, warmup, budget_s=240.0):
    """`steps` identical steps after `warmup` (at most one: a CPU has no clocks to ramp); on a host too slow for that
    within `budget_s` the number of timed steps is cut (never the step itself) and reported as steps_run."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cases, nbytes, sample = cpu_cases()
    t = time.perf_counter()
    for _ in range(min(warmup, 1)):
        cpu_step(cases)
    est = time.perf_counter() - t
    if est > 0:
        steps = max(1, min(steps, int(budget_s / est)))
    t = time.perf_counter()
    for _ in range(steps):
        cpu_step(cases)
    dt = (time.perf_counter() - t) / steps
    return dict(value=nbytes / dt / 1e9, ms=dt * 1e3, cores=torch.get_num_threads(), sample=sample, steps_run=steps)


def reduce_max(values, device, world):
    """MAX over ranks of a list of floats (the contract's 'max over ranks' timing)."""
    if world <= 1:
        return [float(v) for v in values]
    t = torch.tensor(values, device=device, dtype=torch.float64)
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    return [float(x) for x in t.tolist()]


def bind_to_gpu_numa_node(index):
    """Pin this process to the CPU cores next to its GPU before any pinned host buffer is allocated, so
    the e2e leg's host buffers sit on the GPU's NUMA node (with N ranks on one box, buffers on the wrong
    socket turn every copy into inter-socket traffic).  Returns what was done, for the JSON line."""
    try:
        import pynvml
        pynvml.nvmlInit()
        p = torch.cuda.get_device_properties(index)     # NVML order != CUDA order under CUDA_VISIBLE_DEVICES
        bus_id = "%08x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bus_id.encode())
        pynvml.nvmlDeviceSetCpuAffinity(h)
        return "cpus near gpu %d (%d cores)" % (index, len(os.sched_getaffinity(0)))
    except Exception as e:      # no NVML / not permitted: keep the default placement
        return "default (%s)" % type(e).__name__


# ---------------------------------------------------------------------------------------------
# Callers: the other half of BASELINE.json's metric ("KineT frames/sec at 1/2/4/8 B200") and configs[3]
# ---------------------------------------------------------------------------------------------
def measure_callers(rank, world, dev, frames=120, train_steps=10):
    """The callers of the op, through the product API, device-timed, max over ranks:

    tracking frames/s (BASELINE.json configs[2]) -- kinet_b200.tracking.SequenceTracker: the reference's per-frame
      loop (tracker.py:269-557) over kinet_b200.modules.DeformableTransformer + FrameGraphs on a synthetic MOT17
      sequence (750x1333 pyramid, d=288, 6+6 layers, 500 object queries + the previous frame's surviving tracks as
      track queries, so the query count changes every frame), bf16 autocast, one read-back per frame; one
      independent sequence per GPU (replicas);
    training samples/s (configs[3]) -- kinet_b200.training.GraphedTrainStep over the same transformer class
      (800x1333 pyramid, d=256, batch 2 per GPU, dropout 0.1, the reference's optimizer groups and gradient clipping,
      src/train.py:103-117, engine.py:147): fwd+bwd and AdamW as two CUDA graphs around ONE NCCL all-reduce.

    The backbone, the matcher / losses and the datasets are the reference's and are not part of these numbers."""
    from torch import nn
    from kinet_b200.modules import DeformableTransformer
    from kinet_b200.tracking import DetectionHeads, SequenceTracker
    from kinet_b200.training import GraphedTrainStep, reference_param_groups
    from kinet_b200.workloads import PYRAMIDS
    out = {}
    # ---- tracking ----
    levels, d, nq = PYRAMIDS["mot17_750x1333"], 288, 500
    torch.manual_seed(1000 + rank)
    tr = DeformableTransformer(d, 8, 6, 6, 1024, 0.1, "relu", True, len(levels), 4, 4).to(dev).eval()
    heads = DetectionHeads(d, 1).to(dev).eval()
    with torch.no_grad():                       # random weights: let detections come and go around the threshold
        heads.class_embed.bias.zero_()
        nn.init.normal_(heads.class_embed.weight, std=0.5)
        nn.init.normal_(heads.bbox_embed.layers[-1].weight, std=0.02)
    qe = torch.randn(nq, 2 * d, device=dev)
    g = torch.Generator(device=dev).manual_seed(7 + rank)
    scene = [torch.randn(1, d, h, w, device=dev, generator=g) for h, w in levels]
    pos = [torch.randn(1, d, h, w, device=dev, generator=g) * 0.1 for h, w in levels]
    masks = [torch.zeros(1, h, w, dtype=torch.bool, device=dev) for h, w in levels]
    feeds = [[s_ + 0.3 * torch.randn(s_.shape, device=dev, generator=g) for s_ in scene] for _ in range(8)]
    tracker = SequenceTracker(tr, heads, qe, img_size=(1080, 1920), use_graphs=True, bucket=32,
                              autocast_dtype=torch.bfloat16)
    # calibrate the score threshold on the first frame so that ~8 % of the object queries fire (T ~ 20-60 like MOT17)
    tracker.det_thresh = tracker.trk_thresh = 2.0
    tracker.step(feeds[0], masks, pos)
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        hs, _, init_ref, inter_ref, _, _ = tr(feeds[0], masks, pos, qe, None)
        sc = heads(hs[-1], inter_ref[-2])[0].sigmoid().max(-1)[0].flatten()
    thr = float(torch.quantile(sc.float(), 0.92))
    tracker.det_thresh = tracker.trk_thresh = thr
    tracker.reset()
    counts = []
    for i in range(24):                         # warm-up: records the decoder graphs of the buckets that occur
        tracker.step(feeds[i & 7], masks, pos)
    torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in tracker.run((feeds[i & 7], masks, pos) for i in range(frames)):
        counts.append(len(tracker.track_ids) + nq)
    b.record()
    torch.cuda.synchronize()
    ms = reduce_max([a.elapsed_time(b)], dev, world)[0]
    out["tracking_frames_per_s"] = {
        "value": world * frames / ms * 1e3, "ms_per_frame_per_gpu": ms / frames,
        "queries_min_max": [min(counts), max(counts)], "tracks_started": tracker.next_id,
        "decoder_graph_buckets": tracker._graphs.recorded_buckets,
        "config": "kinet_b200.tracking.SequenceTracker: MOT17 750x1333 pyramid, S=20906, d=288 (D=36), 6 enc + 6 dec "
                  "layers, 500 object + T track queries (T changes every frame), bf16 autocast, FrameGraphs, one "
                  "read-back per frame, the next frame's encoder enqueued before the host waits for it "
                  "(SequenceTracker.run); one sequence per GPU"}
    del tracker, tr
    # ---- data-parallel training ----
    levels, d = PYRAMIDS["detr_800x1333"], 256

    class Net(nn.Module):
        def __init__(self):
            super().__init__()
            self.transformer = DeformableTransformer(d, 8, 6, 6, 1024, 0.1, "relu", False, len(levels), 4, 4)
            self.query_embed = nn.Embedding(300, 2 * d)

        def forward(self, srcs, pos):
            if getattr(self, "_masks", None) is None:      # equally sized images: no padding, the same mask tensors every step
                self._masks = [torch.zeros(s_.shape[0], s_.shape[2], s_.shape[3], dtype=torch.bool, device=s_.device) for s_ in srcs]
            return self.transformer(srcs, self._masks, pos, self.query_embed.weight, None)[0]

    torch.manual_seed(0)                                          # same initial weights on every rank
    model = Net().to(dev).train()
    g = torch.Generator(device=dev).manual_seed(100 + rank)       # a different batch on every rank
    L = len(levels)

    def batch():
        srcs = [torch.randn(2, d, h, w, device=dev, generator=g) for h, w in levels]
        pos_ = [torch.randn(2, d, h, w, device=dev, generator=g) * 0.1 for h, w in levels]
        return tuple(srcs + pos_ + [torch.randn(2, 300, d, device=dev, generator=g)])
    batches = [batch() for _ in range(2)]

    def loss_fn(*t):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            hs = model(list(t[:L]), list(t[L:2 * L]))
        return (hs.float() - t[-1]).square().mean()
    step = GraphedTrainStep(model, loss_fn,
                            lambda ps: torch.optim.AdamW(ps, lr=2e-4, weight_decay=1e-4, fused=True, capturable=True),
                            batches[0], param_groups=reference_param_groups(model), clip_max_norm=0.1)
    for i in range(3):
        step(*batches[i & 1])
    torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(train_steps):
        step(*batches[i & 1])
    b.record()
    torch.cuda.synchronize()
    ms = reduce_max([a.elapsed_time(b)], dev, world)[0]
    out["training_samples_per_s"] = {
        "value": world * 2 * train_steps / ms * 1e3, "ms_per_step": ms / train_steps,
        "config": "kinet_b200.modules.DeformableTransformer + GraphedTrainStep: 800x1333 pyramid, S=22223, d=256, "
                  "6 enc + 6 dec layers, 300 queries, batch 2 per GPU, dropout 0.1, bf16 autocast, fwd+bwd + gradient "
                  "clipping 0.1 + AdamW with the reference's 3 learning-rate groups; two CUDA graphs + one NCCL "
                  "gradient all-reduce per step; synthetic MSE loss (matcher / set criterion are the reference's)"}
    del step, model
    # ---- the literal `kine` path: detection-token transformer, no MSDeformAttn in it ----
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "tools"))
    import kine_bench
    k = kine_bench.measure(dev, frames=200, seed=rank)
    kt = kine_bench.measure_tracker(dev, frames=200, seed=rank)
    kr = kine_bench.measure_tracker(dev, frames=200, seed=rank, raw=True)
    ktr = kine_bench.measure_training(dev, steps=20)
    (kms,) = reduce_max([ktr["train_graph_ms_per_step"]], dev, world)
    ms, ms_tr = reduce_max([kt["tracker_ms_per_frame"], k["fused_graphs_ms_per_frame"]], dev, world)
    out["kine_frames_per_s"] = {
        "value": world * 1e3 / ms, "ms_per_frame_per_gpu": ms, "tracks_min_max": kt["tracks_min_max"],
        "tracks_started": kt["tracks_started"], "graphs_recorded": kt["graphs_recorded"],
        "from_raw_detections": {"ms_per_frame_per_gpu": kr["tracker_ms_per_frame"], "tracks_min_max": kr["tracks_min_max"],
                                "host_enqueue_ms": kr["phases_ms"]["enqueue"],
                                "note": "the same loop fed raw boxes + confidences, the model's own backbones "
                                        "(modules.kinematic_backbone) recorded in the graph: one pinned block + one replay per "
                                        "frame; with random weights only a few tracks live at a time (a lighter host update)"},
        "training": {"samples_per_s": world * ktr["batch"] * 1e3 / kms, "ms_per_step_per_gpu": kms, "batch_per_gpu": ktr["batch"],
                     "same_parameters_through_stock_pytorch_modules_ms": {"eager": ktr["train_plain_eager_ms"],
                                                                          "one_cuda_graph": ktr["train_plain_graph_ms"]},
                     "note": "cfgs/train_kinet.yaml shape (batch 8 x 150 tokens, 150 queries, d 288, 1 + 1 layers, dropout 0.1), "
                             "synthetic loss, clip 0.1 + AdamW inside kinet_b200.training.GraphedTrainStep (data-parallel: "
                             "one gradient all-reduce per step when N > 1)"},
        "transformer_only": {"frames_per_s": world * 1e3 / ms_tr, "ms_per_frame_per_gpu": ms_tr,
                             "same_parameters_through_stock_pytorch_modules_ms": {"eager": k["plain_eager_ms"],
                                                                                  "one_cuda_graph": k["plain_graph_ms"]}},
        "config": "kinet_b200.tracking.KineSequenceTracker (the loop of TrackerKinematic.step: trails -> track queries, "
                  "detector replay from KineFrameGraphs, one read-back, host thresholds / NMS / births / terminations) over "
                  "kinet_b200.modules.kinematic_detector.KinematicDetectorTransformer: " + k["config"] + ", " + k["dtype"] +
                  "; transformer_only: the detector's transformer + heads alone with token and track-query counts that "
                  "change every frame; one sequence per GPU"}
    return out


def measure_sweep(dev, iters=5):
    """BASELINE.json configs[4] (MOT20 crowd stress): full-res 1080x1920 pyramid, encoder form and the decoder form
    at 300 object + 500 track queries, forward and backward, bf16 and fp32 values; CUDA events, L2 flushed between
    iterations (a 256 MB buffer is overwritten)."""
    from kinet_b200.functions import ms_deform_attn_backward, ms_deform_attn_forward
    peak, _ = measured_peak()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    out = {}
    for tag in ("C5_enc", "C5_dec"):
        for name, dt in (("bf16", torch.bfloat16), ("f32", torch.float32)):
            s = resolve(tag)
            value, shapes, loc, attn = make_inputs(tag, dev, dt, torch.float32, "structured", 3)
            go = torch.randn(s["N"], s["Lq"], s["M"] * s["D"], device=dev).to(dt)
            bf, bb = algorithmic_bytes(s, value.element_size(), 4)
            res = {}
            for what, fn, nbytes in (("fwd", lambda: ms_deform_attn_forward(value, shapes, loc, attn, 64), bf),
                                     ("bwd", lambda: ms_deform_attn_backward(value, shapes, loc, attn, go, 64), bb)):
                for _ in range(2):
                    fn()
                ts = []
                for _ in range(iters):
                    flush.zero_()
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    torch.cuda._sleep(200000)
                    a.record(); fn(); b.record()
                    torch.cuda.synchronize()
                    ts.append(a.elapsed_time(b) * 1e3)
                ts.sort()
                us = ts[len(ts) // 2]
                res[what] = {"us": us, "GBps": nbytes / us / 1e3, "frac_of_hbm": nbytes / us / 1e3 / peak}
            out["%s_%s" % (tag, name)] = res
    out["config"] = ("MOT20 full-res pyramid (135,240)(68,120)(34,60)(17,30), S=43110, N=1, M=8, D=32, L=4, P=4; "
                     "C5_enc: Lq=S; C5_dec: Lq=800 (300 object + 500 track queries); backward incl. zero-fill / rounding")
    return out


def copy_ceiling(host, sink, dev, world, steps=3):
    """What the box gives for the e2e leg's transfers alone: the same pinned host tensors copied H2D and the same
    result sizes copied D2H, concurrently on two streams, one cudaMemcpyAsync per tensor, no kernels."""
    land = [tuple(torch.empty_like(t, device=dev) for t in tup) for tup in host]
    res = [tuple(torch.empty_like(t, device=dev) for t in tup) for tup in sink]
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()

    def once():
        with torch.cuda.stream(s_in):
            for hs_, ds_ in zip(host, land):
                for h_, d_ in zip(hs_, ds_):
                    d_.copy_(h_, non_blocking=True)
        with torch.cuda.stream(s_out):
            for hs_, ds_ in zip(sink, res):
                for h_, d_ in zip(hs_, ds_):
                    h_.copy_(d_, non_blocking=True)
    once()
    torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    cur = torch.cuda.current_stream()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    s_in.wait_event(a); s_out.wait_event(a)
    for _ in range(steps):
        once()
    cur.wait_stream(s_in); cur.wait_stream(s_out)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--e2e-steps", type=int, default=0, help="cap on the e2e leg's timed steps; 0 = --steps")
    ap.add_argument("--cpu-steps", type=int, default=3)
    ap.add_argument("--no-sweep", action="store_true", help="skip the configs[4] (MOT20 crowd) per-call sweep")
    ap.add_argument("--no-callers", action="store_true",
                    help="skip the frames/s and training samples/s leg (tracking / training stacks around the op)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    config = {"workload": WORKLOAD, "pyramid": "800x1333 -> (100,167)(50,84)(25,42)(13,21)",
              "calls_per_step": "6 encoder-form (N=2,Lq=S=22223) + 6 decoder-form (Lq=300), fwd+bwd each",
              "M": 8, "D": 32, "L": 4, "P": 4, "value_dtype": "bf16", "loc_attn_dtype": "f32",
              "l2": "every layer has its own inputs: ~0.7 GB touched per step > 126 MB L2 (no flush needed)",
              "parallelism": "replicas x%d (independent batches per GPU, no collective on the path)" % world}

    if args.impl == "reference":
        if rank != 0:
            return
        r = run_cpu(args.steps, args.warmup)
        line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms"], "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
                "steps_run": r["steps_run"],
                "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port",
                                 "sample": r["sample"]},
                "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line), flush=True)
        return

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the MSDeformAttn path has no CPU fallback "
                         "(use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # N = 1 keeps every core: the cpu_baseline leg of that run uses all host threads
    config["host_affinity"] = bind_to_gpu_numa_node(local) if world > 1 else "all cores (single rank)"
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=dev)
    r = run_ours(args, rank, world, dev)
    ms, e2e_ms = r["ms"], r["e2e_ms"]
    ms, e2e_ms, ceiling_ms = reduce_max([ms, e2e_ms, r["ceiling_ms"]], dev, world)
    sweep = None
    if rank == 0 and world == 1 and not args.no_sweep:
        try:
            sweep = measure_sweep(dev)
        except Exception as e:
            sweep = {"error": "%s: %s" % (type(e).__name__, str(e)[:200])}
    callers = None
    if not args.no_callers:
        try:
            callers = measure_callers(rank, world, dev)
        except Exception as e:          # reported, never fatal for the contract line
            callers = {"error": "%s: %s" % (type(e).__name__, str(e)[:200])}
    if rank == 0:
        peak, peak_src = measured_peak()
        ms_step = ms / args.steps
        value = world * r["total_bytes"] / (ms_step * 1e-3) / 1e9
        e2e = world * r["total_bytes"] / (e2e_ms * 1e-3) / 1e9
        kms = sorted(r["kern_ms"])
        kavg = sum(kms) / len(kms)
        achieved = r["enc_bwd_bytes"] / (kavg * 1e-3) / 1e9
        cpu = run_cpu(args.cpu_steps, 1)
        ri = roofline_inputs()
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": config,
                "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"],
                        "ms_per_step": e2e_ms, "steps": r["e2e_steps"], "host_enqueue_ms_per_step": r["e2e_host_ms"],
                        "copy_ceiling_ms_per_step": ceiling_ms, "frac_of_copy_ceiling": ceiling_ms / e2e_ms,
                        "copy_ceiling": "the same pinned tensors copied H2D and the same result sizes copied D2H "
                                        "concurrently on two streams, one cudaMemcpyAsync per tensor, no kernels, "
                                        "all ranks at once (max over ranks)",
                        "api": "kinet_b200.MSDeformAttnFunction.apply + Tensor.backward, pinned host tensors; H2D, kernels and D2H on three streams"},
                "gpu_launches": r["launches"],
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                             "frac": achieved / peak, "traffic": ri.get("traffic"), "traffic_source": ri.get("source"),
                             "traffic_stale": ri.get("stale"), "peak_source": peak_src,
                             "kernel": "encoder-form backward (N=2, Lq=S=22223, bf16 value / f32 loc+attn): "
                                       "dense::msda_backward_dense_kernel<bf16,f32,D=32> (finest level's query tiles, "
                                       "tcgen05 + TMEM) + msda_backward_kernel<bf16,f32,VEC=4,GS=8,NW=8> in split mode "
                                       "(the coarser levels' queries); the two launches are timed as one",
                             "kernel_us_avg": kavg * 1e3, "kernel_us_min": kms[0] * 1e3,
                             "launches_timed": len(kms), "algorithmic_bytes_per_launch": r["enc_bwd_bytes"],
                             "step_frac_of_hbm": value / world / peak,
                             "binding_limit": {
                                 "what": "not HBM (dram throughput < 10 %): the dense kernel is bound by the SM's L1 / "
                                         "shared-memory data pipe and instruction issue (per-corner scalar work of one "
                                         "thread per query row), the per-row kernel by the L2 fp32 reduction rate",
                                 "red_sectors_per_launch": ri.get("red_sectors"),
                                 "red_sectors_per_launch_round1": 80064552,
                                 "dense_kernel_us_ncu": ri.get("dense_us"), "per_row_kernel_us_ncu": ri.get("rows_us"),
                                 "l1_data_pipe_pct_dense": ri.get("l1_data_pipe_pct"),
                                 "source": ri.get("source")},
                             "note": "see DESIGN.md 5.2 / 5.5"},
                "cpu_baseline": {"value": cpu["value"], "unit": UNIT, "cores": cpu["cores"], "kind": "port",
                                 "sample": cpu["sample"], "ms_per_sample": cpu["ms"]},
                "clocks": r["clocks"]}
        if callers is not None:
            line["callers"] = callers
        if sweep is not None:
            line["sweep"] = sweep
        print(json.dumps(line), flush=True)
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
