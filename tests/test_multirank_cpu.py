"""world_size-2 gloo tests (CPU) of the N>1 host logic: max-over-ranks timing reduction used by
bench.py, the --impl reference rank-0-only rule, and round-robin sharding of independent units."""
import json
import os
import subprocess
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    import bench
    from kinet_b200.sharding import shard_round_robin
    dist.init_process_group("gloo", rank=rank, world_size=world)
    got = bench.reduce_max([10.0 + rank, 5.0 - rank], torch.device("cpu"), world)
    mine = shard_round_robin(list(range(7)), rank, world)
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    q.put((rank, got, gathered))
    dist.destroy_process_group()


def test_reduce_max_and_sharding_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    from kinet_b200.sharding import merge_round_robin
    for rank, got, gathered in res:
        assert got == [11.0, 5.0]
        assert merge_round_robin(gathered) == list(range(7))


def test_sharding_edges():
    from kinet_b200.sharding import merge_round_robin, shard_round_robin
    assert shard_round_robin([], 0, 4) == []
    assert shard_round_robin([1, 2, 3], 2, 8) == [3]
    assert shard_round_robin([1, 2, 3], 5, 8) == []
    assert merge_round_robin([[0, 2], [1]]) == [0, 1, 2]


def test_reference_arm_only_rank0_prints():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                        "--steps", "1", "--warmup", "1"], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_reference_arm_json_line():
    env = dict(os.environ, RANK="0", WORLD_SIZE="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                        "--warmup", "1"], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "GB/s" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["config"]["workload"] == "detr_6enc_6dec_bf16"


def _train_worker(rank, world, port, q):
    """Data-parallel step the way tools/train_bench.py does it: gradients of every parameter land in
    FlatParams.grads, ONE all-reduce of that buffer, one optimizer step on the flat parameter."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    from kinet_b200.training import FlatParams
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)                                   # same initial weights on every rank
    model = torch.nn.Sequential(torch.nn.Linear(12, 20), torch.nn.ReLU(), torch.nn.Linear(20, 5), torch.nn.LayerNorm(5))
    fp = FlatParams(model)
    opt = torch.optim.AdamW([fp.flat_parameter], lr=1e-2, weight_decay=1e-2)
    g = torch.Generator().manual_seed(7)
    x, y = torch.randn(8, 12, generator=g), torch.randn(8, 5, generator=g)
    lo, hi = rank * 4, rank * 4 + 4                       # this rank's shard of the global batch
    for _ in range(2):
        fp.zero_grad()
        (model(x[lo:hi]) - y[lo:hi]).square().mean().backward()
        dist.all_reduce(fp.grads, op=dist.ReduceOp.SUM)   # gloo has no AVG
        fp.grads.div_(world)
        opt.step()
        fp.refresh()
    q.put((rank, fp.master.numpy().copy(), fp.lp.float().numpy().copy()))     # numpy: pickled by value
    dist.destroy_process_group()


def test_flat_params_data_parallel_world2_matches_single_process():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_train_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted((q.get(timeout=120) for _ in procs), key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    res = [(r, torch.from_numpy(m), torch.from_numpy(l)) for r, m, l in res]
    assert torch.equal(res[0][1], res[1][1])              # replicas stay identical
    # the same two steps in one process on the whole batch (mean over 8 = average of the two shard means)
    from kinet_b200.training import FlatParams
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(12, 20), torch.nn.ReLU(), torch.nn.Linear(20, 5), torch.nn.LayerNorm(5))
    plain = [p.detach().clone() for p in model.parameters()]
    opt = torch.optim.AdamW(model.parameters(), lr=1e-2, weight_decay=1e-2)
    g = torch.Generator().manual_seed(7)
    x, y = torch.randn(8, 12, generator=g), torch.randn(8, 5, generator=g)
    for _ in range(2):
        opt.zero_grad()
        (model(x) - y).square().mean().backward()
        opt.step()
    fp = FlatParams(torch.nn.Sequential(torch.nn.Linear(12, 20), torch.nn.ReLU(), torch.nn.Linear(20, 5),
                                        torch.nn.LayerNorm(5)))
    # walk the flat master buffer of rank 0 with FlatParams' own layout (64-element aligned slots)
    off = 0
    for p, p0 in zip(model.parameters(), plain):
        n = p.numel()
        got = res[0][1][off:off + n].view_as(p)
        assert not torch.equal(got, p0)                   # the parameters did move
        assert torch.allclose(got, p.detach(), rtol=1e-5, atol=1e-6)
        assert torch.allclose(res[0][2][off:off + n].view_as(p), p.detach().bfloat16().float(), rtol=1e-2, atol=1e-2)
        off += -(-n // 64) * 64
    assert off == fp.master.numel()
