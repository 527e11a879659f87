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


def _reducer_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from fod_b200.ddp import BucketedGradReducer
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.ReLU(), torch.nn.Linear(5, 3), torch.nn.Linear(3, 1))
    net[2].bias.requires_grad_(False)  # frozen parameters take no slot
    red = BucketedGradReducer(net.parameters(), bucket_mb=64 / (1 << 20))  # 16-float buckets -> several buckets
    res = {"n_buckets": len(red.buckets)}
    for it in range(2):  # second iteration: views survive zero_grad, counters re-arm
        red.zero_grad()
        x = torch.arange(12, dtype=torch.float32).view(2, 6) * 0.1 + rank + it
        net(x).sum().backward()
        red.finish()
        res["grads%d" % it] = [p.grad.clone() for p in net.parameters() if p.requires_grad]
    res["norm"] = float(red.clip_grad_norm_(0.1))
    res["clipped_norm"] = float(torch.cat([p.grad.flatten() for p in net.parameters() if p.requires_grad]).norm())
    try:  # replacing a view must be caught, not silently drop the parameter out of the all-reduce
        net[0].weight.grad = None
        red.zero_grad()
        net(torch.ones(1, 6)).sum().backward()
        res["caught"] = False
    except (RuntimeError, AttributeError):
        res["caught"] = True
    out[rank] = res
    dist.destroy_process_group()


def test_bucketed_grad_reducer_two_ranks():
    """fod_b200.ddp.BucketedGradReducer == DDP semantics (mean of per-rank gradients), multi-bucket, re-armed per step,
    flat clipping == torch.nn.utils.clip_grad_norm_."""
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_reducer_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    assert out[0]["n_buckets"] >= 3
    for it in range(2):
        torch.manual_seed(0)
        net = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.ReLU(), torch.nn.Linear(5, 3), torch.nn.Linear(3, 1))
        net[2].bias.requires_grad_(False)
        tot = 0
        for r in range(world):
            x = torch.arange(12, dtype=torch.float32).view(2, 6) * 0.1 + r + it
            tot = tot + net(x).sum() / world
        tot.backward()
        expect = [p.grad for p in net.parameters() if p.requires_grad]
        for r in range(world):
            for g, e in zip(out[r]["grads%d" % it], expect):
                assert torch.allclose(g, e, rtol=1e-5, atol=1e-6)
    ref_norm = float(torch.nn.utils.clip_grad_norm_([p for p in net.parameters() if p.requires_grad], 0.1))
    assert abs(out[0]["norm"] - ref_norm) < 1e-4 * ref_norm
    assert abs(out[0]["clipped_norm"] - min(0.1, ref_norm)) < 1e-5
    assert out[0]["caught"] and out[1]["caught"]


def test_flat_train_state_layout_cpu():
    """fod_b200.optim.FlatTrainState on CPU (no kernel call): parameters / gradients become views of the arenas with
    16-byte slots, the backbone group sits last (lower LR, runs/_helper.py:93-100), names and values are unchanged."""
    from fod_b200.optim import FlatTrainState
    torch.manual_seed(0)
    net = torch.nn.ModuleDict({"backbone": torch.nn.Linear(5, 3), "head": torch.nn.Linear(3, 7)})
    before = {k: v.clone() for k, v in net.state_dict().items()}
    flat = FlatTrainState(net, lr=1e-4, lr_backbone=1e-5, bucket_mb=1e-5)
    assert list(net.state_dict()) == list(before) and all(torch.equal(net.state_dict()[k], before[k]) for k in before)
    offs = sorted(flat.reducer.offset.values())
    assert all(o % 4 == 0 for o in offs) and flat.n % 4 == 0
    back = [flat.reducer.offset[p] for n, p in net.named_parameters() if "backbone" in n]
    head = [flat.reducer.offset[p] for n, p in net.named_parameters() if "backbone" not in n]
    assert max(head) < min(back) == flat.group_begin[1] and flat.group_lr == [1e-4, 1e-5]
    net["head"](net["backbone"](torch.ones(2, 5))).sum().backward()
    w = net["head"].weight
    o = flat.reducer.offset[w]
    assert torch.equal(flat.grads[o:o + w.numel()].view_as(w), w.grad) and float(w.grad.abs().sum()) > 0
    assert torch.equal(flat.params[o:o + w.numel()].view_as(w), w.data)


def test_fused_param_groups_get_adjacent_slots_cpu():
    """Modules that read several parameters as one concatenated operand announce them through fused_param_groups():
    FlatTrainState gives them adjacent arena slots in that order, autograd.FusedParams exposes data / grad / bf16 copy as
    one strided view each, and the reducer's ready signal of the view counts every member's bucket down."""
    from fod_b200 import autograd as A
    from fod_b200.optim import FlatTrainState

    class Proj(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.a, self.other, self.b, self.c = (torch.nn.Linear(8, 8) for _ in range(4))

        def fused_param_groups(self):
            return [self.c.weight, self.a.weight, self.b.weight], [self.c.bias, self.a.bias, self.b.bias]

    torch.manual_seed(0)
    net = Proj()
    ref = torch.cat([net.c.weight, net.a.weight, net.b.weight]).detach().clone()
    flat = FlatTrainState(net, bucket_mb=1e-4)   # several buckets: the group straddles bucket boundaries
    assert len(flat.reducer.buckets) > 1
    gw, gb = net.fused_param_groups()
    fw, fb = A.FusedParams.build(gw), A.FusedParams.build(gb)
    assert fw is not None and fb is not None and fw.shape == (24, 8) and fb.shape == (24,)
    assert torch.equal(fw.data, ref) and torch.equal(fw._fod_shadow.float(), flat.shadow[flat.reducer.offset[net.c.weight]:][:192].view(24, 8).float())
    fw.grad.fill_(2.0)
    assert all(float(p.grad.min()) == 2.0 for p in gw) and float(net.other.weight.grad.abs().sum()) == 0.0
    assert A.FusedParams.build([net.a.weight, net.c.weight]) is None      # not adjacent in this order
    assert A.FusedParams.build(gw, fw) is fw                              # cached view is reused
    with torch.no_grad():
        assert A.FusedParams.build(gw) is None                           # no gradient recording: caller falls back
    flat.begin_step()
    fw._fod_uses = 1
    A._done(fw)
    assert all(p in flat.reducer._ready for p in gw)


def _broadcast_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from fod_b200.optim import FlatTrainState
    torch.manual_seed(100 + rank)  # every process initialises from its OWN RNG, as runs/_model.py does
    net = torch.nn.ModuleDict({"backbone": torch.nn.Sequential(torch.nn.Linear(5, 3), torch.nn.BatchNorm1d(3)),
                               "head": torch.nn.Linear(3, 7)})
    net["backbone"][0].bias.requires_grad_(False)          # frozen parameter: outside the arena
    net["backbone"][1].running_mean.fill_(float(rank + 1))  # buffer
    FlatTrainState(net, bucket_mb=1e-5)
    out[rank] = {k: v.clone() for k, v in net.state_dict().items()}
    dist.destroy_process_group()


def test_flat_train_state_broadcasts_rank0_like_ddp():
    """runs/_model.py:76-82: DistributedDataParallel's constructor makes every replica start from rank 0's parameters
    and buffers.  FlatTrainState replaces that wrapper and must do the same (trainable arena, frozen parameters,
    buffers), or replicas initialised from different RNG streams would silently train different models."""
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_broadcast_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    torch.manual_seed(100)
    ref = torch.nn.ModuleDict({"backbone": torch.nn.Sequential(torch.nn.Linear(5, 3), torch.nn.BatchNorm1d(3)),
                               "head": torch.nn.Linear(3, 7)})
    ref["backbone"][1].running_mean.fill_(1.0)
    for k, v in ref.state_dict().items():
        assert torch.equal(out[0][k], v) and torch.equal(out[1][k], v), k


def test_flat_train_state_checkpoint_is_torch_adamw_format():
    """FlatTrainState.state_dict() is torch.optim.AdamW's format in the reference's group order (runs/_helper.py:84-107),
    keyed by name as well: a reference checkpoint's `optimizer` entry (trainer.py:282-299) resumes here, a changed arena
    layout cannot misplace moments, and a mismatched entry is refused."""
    import pytest
    from fod_b200.optim import FlatTrainState

    def make():
        torch.manual_seed(0)
        return torch.nn.ModuleDict({"backbone": torch.nn.Linear(5, 3), "head": torch.nn.Linear(3, 7), "z": torch.nn.Linear(7, 2)})

    ref = make()
    named = list(ref.named_parameters())
    opt = torch.optim.AdamW([{"params": [p for n, p in named if "backbone" not in n]},
                             {"params": [p for n, p in named if "backbone" in n], "lr": 1e-5}], lr=1e-4, weight_decay=1e-4)
    for _ in range(3):
        opt.zero_grad()
        ref["z"](ref["head"](ref["backbone"](torch.ones(2, 5)))).sum().backward()
        opt.step()
    sd_ref = opt.state_dict()
    net = make()
    flat = FlatTrainState(net, lr=3e-4, lr_backbone=3e-5, bucket_mb=1e-5)
    flat.load_state_dict(sd_ref)  # torch's own format, no names: the reference's group order
    assert float(flat.state[2]) == 3.0 and flat.group_lr == [1e-4, 1e-5]
    idx = 0
    for grp in (False, True):
        for n, p in net.named_parameters():
            if ("backbone" in n) == grp:
                o = flat.reducer.offset[p]
                assert torch.equal(flat.exp_avg[o:o + p.numel()].view_as(p), sd_ref["state"][idx]["exp_avg"]), n
                assert torch.equal(flat.exp_avg_sq[o:o + p.numel()].view_as(p), sd_ref["state"][idx]["exp_avg_sq"]), n
                idx += 1
    mine = flat.state_dict()
    assert mine["param_names"][-2:] == ["backbone.weight", "backbone.bias"] and len(mine["state"]) == 6
    fresh = torch.optim.AdamW([{"params": [p for n, p in named if "backbone" not in n]},
                               {"params": [p for n, p in named if "backbone" in n]}])
    fresh.load_state_dict({k: v for k, v in mine.items() if k != "param_names"})  # torch accepts it back
    assert torch.equal(fresh.state_dict()["state"][4]["exp_avg"], sd_ref["state"][4]["exp_avg"])
    net2 = make()
    flat2 = FlatTrainState(net2, bucket_mb=64.0)  # another bucket layout: names, not offsets, place the moments
    flat2.load_state_dict(mine)
    w = net2["z"].weight
    o = flat2.reducer.offset[w]
    assert torch.equal(flat2.exp_avg[o:o + w.numel()].view_as(w), sd_ref["state"][2]["exp_avg"])
    bad = {**mine, "state": {**mine["state"], 0: {**mine["state"][0], "exp_avg": torch.zeros(3, 3)}}}
    with pytest.raises(ValueError):
        flat2.load_state_dict(bad)
