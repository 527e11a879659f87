"""World-size-2 gloo tests (CPU) of the N>1 host logic: num_boxes exchange (set_criterion.py:185-193), max-over-ranks
timing, stat averaging (utils/distributed.py:63-75), per-rank sharding seeds, and DDP gradient averaging semantics that
bench.py --gpus N relies on (weak scaling: each rank owns its clips)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from fod_b200 import dist as D
    res = {}
    local = torch.tensor([float([0, 37][rank])])
    res["num_boxes"] = float(D.global_num_boxes(local))                   # (0 + 37) / 2
    res["num_boxes_empty"] = float(D.global_num_boxes(torch.zeros(1)))      # clamp to 1
    res["num_boxes_local"] = float(D.global_num_boxes(local, distributed=False))
    res["max_ms"] = D.max_over_ranks([12.5, 40.25][rank])
    st = D.reduce_stats({"labels": torch.tensor(float(rank)), "box_l1": torch.tensor(2.0 + rank)})
    res["stats"] = {k: float(v) for k, v in st.items()}
    res["seeds"] = [D.shard_seed(rank, i) for i in range(3)]
    # DDP averaging: each rank's loss is a sum over ITS clips; the averaged gradient equals the single-process gradient
    # of the mean over ranks — the reference's semantics (runs/_model.py:76-82).
    torch.manual_seed(0)
    lin = torch.nn.Linear(4, 1)
    ddp = torch.nn.parallel.DistributedDataParallel(lin)
    x = torch.arange(8, dtype=torch.float32).view(2, 4) + 10 * rank
    ddp(x).sum().backward()
    res["grad"] = lin.weight.grad.clone()
    out[rank] = res
    dist.destroy_process_group()


def test_two_rank_gloo_plumbing():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    r0, r1 = out[0], out[1]
    assert r0["num_boxes"] == r1["num_boxes"] == 18.5
    assert r0["num_boxes_empty"] == 1.0 and r0["num_boxes_local"] == 1.0 and r1["num_boxes_local"] == 37.0
    assert r0["max_ms"] == r1["max_ms"] == 40.25
    assert r0["stats"] == r1["stats"] == {"box_l1": 2.5, "labels": 0.5}
    assert not set(r0["seeds"]) & set(r1["seeds"])
    xs = [torch.arange(8, dtype=torch.float32).view(2, 4) + 10 * r for r in range(2)]
    expect = sum(x.sum(0, keepdim=True) for x in xs) / 2
    assert torch.allclose(r0["grad"], expect) and torch.allclose(r1["grad"], expect)
