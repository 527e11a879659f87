"""CUDA path of the epoch-level AP bookkeeping (fod_b200.od_map): pinned-slab side-stream copies with
no host sync per step, aggregation of all thresholds on the device; against the reference golden."""
import pytest
import torch

from conftest import load_golden
from test_od_map_cpu import _same, _trainer_concat

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_accumulator_and_aggregation_on_the_device():
    from fod_b200.od_map import ApAccumulator, aggregate_mean_average_precision
    gold = load_golden("ap_aggregate_4x3")
    acc = ApAccumulator(slab_floats=1 << 15)
    side = torch.cuda.Stream()
    for it in gold["iterations"]:
        with torch.cuda.stream(side):  # the producer is not the default stream: the copy must order after it
            stuffs = [t.to(DEV, non_blocking=True) for t in it["stuffs"]]
            acc.add(stuffs)
    assert all(s.is_pinned() for s in acc._slabs) and all(it[1] is not None for it in acc._items)
    for got, want in zip(acc.tensors(), _trainer_concat(gold)):
        assert got.dtype == want.dtype and torch.equal(got, want)
    for ap in (acc.aggregate(device=DEV), aggregate_mean_average_precision(*_trainer_concat(gold), device=DEV)):
        for k, v in gold["ap"].items():
            assert _same(ap[k], v, 1e-5), k
