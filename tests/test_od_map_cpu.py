"""Epoch-level AP aggregation (future_od/utils/od_map.py:288-364 driven as trainer.py:155,205-236):
the oracle and the drop-in (`fod_b200.od_map`) against a golden from the unmodified reference
(tests/golden/make_golden_ap.py), plus the world-size-2 gloo path of the accumulator."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import GOLDEN, load_golden
from oracle import fod_oracle as O


def _same(a, b, tol=0.0):
    a, b = torch.as_tensor(a).double(), torch.as_tensor(b).double()
    if a.shape != b.shape or not torch.equal(torch.isnan(a), torch.isnan(b)) or not torch.equal(torch.isinf(a), torch.isinf(b)):
        return False
    fin = torch.isfinite(b)
    return bool(((a[fin] - b[fin]).abs() <= tol).all()) and torch.equal(a[~fin & ~torch.isnan(b)], b[~fin & ~torch.isnan(b)])


def _trainer_concat(gold):
    """trainer.py:155,208-211,231-236 for a single process."""
    lst = [[], [], [], []]
    for it in gold["iterations"]:
        for i, t in enumerate(it["stuffs"]):
            lst[i].append(t)
    return torch.cat(lst[0], dim=2), torch.cat(lst[1], dim=2), torch.cat(lst[2], dim=2), torch.stack(lst[3], dim=2)


def test_oracle_matches_reference_golden():
    gold = load_golden("ap_aggregate_4x3")
    for it in gold["iterations"]:  # a second pin of the per-step bookkeeping (od_map.py:214-287) on hit-rich inputs
        st = O.od_map_stuffs(it["pred_boxes"], it["scores"], it["boxes"], it["classes"], it["active"], gold["imsize"])
        assert _same(st[0], it["stuffs"][0], 1e-6)
        assert all(torch.equal(a.long(), b.long()) for a, b in zip(st[1:], it["stuffs"][1:]))
    ap = O.aggregate_ap(*_trainer_concat(gold))
    assert set(ap) == set(gold["ap"])
    for k, v in gold["ap"].items():
        assert _same(ap[k], v), k
    assert torch.isnan(gold["ap"]["all"]).any() and torch.isinf(gold["ap"]["all"]).any()  # the edge cases are in the vector
    assert int(_trainer_concat(gold)[1].sum()) > 1000


def test_drop_in_aggregation_matches_reference_golden():
    from fod_b200.od_map import aggregate_mean_average_precision
    gold = load_golden("ap_aggregate_4x3")
    ap = aggregate_mean_average_precision(*_trainer_concat(gold), device="cpu")
    assert set(ap) == set(gold["ap"])
    for k, v in gold["ap"].items():
        assert _same(ap[k], v, 1e-6), k
    assert torch.is_tensor(ap["all"]) and torch.is_tensor(ap["generic"]) and not torch.is_tensor(ap["classavg"])


def test_accumulator_single_process():
    from fod_b200.od_map import ApAccumulator
    gold = load_golden("ap_aggregate_4x3")
    acc = ApAccumulator(slab_floats=1 << 15)  # small slabs: the epoch spans several
    for it in gold["iterations"]:
        acc.add(it["stuffs"])
    assert len(acc) == 4 and len(acc._slabs) > 1
    for got, want in zip(acc.tensors(), _trainer_concat(gold)):
        assert got.dtype == want.dtype and torch.equal(got, want)
    ap = acc.aggregate()
    for k, v in gold["ap"].items():
        assert _same(ap[k], v, 1e-6), k
    acc.reset()
    assert len(acc) == 0
    acc.add(gold["iterations"][0]["stuffs"])
    assert torch.equal(acc.tensors()[0], gold["iterations"][0]["stuffs"][0])


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from fod_b200.od_map import ApAccumulator
    gold = torch.load(os.path.join(GOLDEN, "ap_aggregate_4x3.pt"), weights_only=False)
    from fod_b200 import dist as D
    acc = ApAccumulator(distributed=True)
    merged_live, separate = [], []
    for step in range(2):  # rank r owns iterations r and r + 2
        stats = {"labels": torch.tensor(1.0 + rank + 10 * step), "box_l1": torch.tensor(0.25 * rank),
                 "box_giou": torch.tensor(3.0), "cardinality": torch.tensor(float(step)), "class_error": torch.tensor(50.0 - rank)}
        acc.add(gold["iterations"][rank + 2 * step]["stuffs"], stats)
        merged_live.append({k: float(v) for k, v in acc.last_stats_device.items()})
        separate.append({k: float(v) for k, v in D.reduce_stats(stats).items()})  # the reference's own all-reduce
    out[rank] = {"tensors": acc.tensors(), "ap": acc.aggregate(), "stats": acc.stats_history(), "live": merged_live,
                 "separate": separate}
    dist.destroy_process_group()


def test_accumulator_two_ranks_gloo():
    """utils/distributed.py:78-89 + trainer.py:205-211: every rank ends up with every rank's objects, step-major with
    the ranks in order inside a step; the AP is the single-process one (it is invariant to the object order)."""
    gold = load_golden("ap_aggregate_4x3")
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    its = gold["iterations"]
    order = [0, 1, 2, 3]  # step 0: ranks 0, 1 -> iterations 0, 1; step 1: iterations 2, 3
    want = [torch.cat([its[i]["stuffs"][0] for i in order], dim=2), torch.cat([its[i]["stuffs"][1] for i in order], dim=2),
            torch.cat([its[i]["stuffs"][2] for i in order], dim=2), torch.stack([its[i]["stuffs"][3] for i in order], dim=2)]
    for r in range(2):
        for got, w in zip(out[r]["tensors"], want):
            assert torch.equal(got, w)
        for k, v in gold["ap"].items():
            assert _same(out[r]["ap"][k], v, 1e-6), k
        # the 5 logged scalars rode in the same packed gather: their mean over ranks equals what the
        # reference's separate all-reduce (utils/distributed.py:63-75) gives, live on the device and in the host history
        assert out[r]["stats"] == out[r]["live"] == out[r]["separate"]
        assert out[r]["stats"][1] == {"labels": 11.5, "box_l1": 0.125, "box_giou": 3.0, "cardinality": 1.0, "class_error": 49.5}
