"""Epoch-level AP bookkeeping: what `future_od/trainer.py:155,197-236` does with the per-step
`od_map_stuffs` — gather over ranks (`utils/distributed.py:78-89`), `.cpu()`, concatenate, and at the end of the epoch
`aggregate_mean_average_precision` (`utils/od_map.py:288-364`) — without ever blocking the training stream:

* the four per-step tensors are packed into ONE float32 buffer (bools and counts are exact in fp32), so the N > 1 path
  costs one all-gather per step instead of four;
* the device->host copy goes into a pinned slab on a side stream that waits on the producing stream; the host only
  synchronises on the recorded events when the epoch's result is asked for;
* the aggregation evaluates all IoU thresholds in one sort + gather + cumsum pass on the device instead of a Python
  loop over thresholds with four host->device copies each (`od_map.py:340-347`).

Device-agnostic (CPU tensors take the same code path minus the streams), so the logic is tested with gloo on CPU.
"""
import warnings

import numpy as np
import torch
import torch.distributed as dist


def _average_precision_all_thresholds(confs, is_positive, size_categories, num_annos):
    """`_get_ap` (od_map.py:288-312) for every threshold at once.  confs, is_positive (T, C, N); size_categories
    (C, S, N) bool; num_annos (C, S, iterations) -> (T, C, S) fp32, with the reference's dtypes at every step (bool
    products, int64 cumulative counts, fp32 precision) so that 0/0 -> NaN and x/0 -> inf come out the same."""
    t, c, n = confs.shape
    s = size_categories.shape[1]
    order = confs.argsort(dim=2, descending=True)
    pos = is_positive.bool().gather(2, order)[:, :, None, :]                                  # (T, C, 1, N)
    cat = size_categories.bool()[None].expand(t, c, s, n).gather(3, order[:, :, None, :].expand(t, c, s, n))
    pos = pos & cat                                                                           # (T, C, S, N)
    precision = pos.cumsum(dim=3) / (cat.cumsum(dim=3) + 1e-5)
    return (precision * pos).sum(dim=3) / num_annos.sum(dim=2)[None]


def aggregate_mean_average_precision(confs, is_positive, size_categories, num_annos, device=None):
    """Drop-in for `future_od.utils.od_map.aggregate_mean_average_precision` (od_map.py:315-364): same arguments, same
    dict — "all" (T, C-1, S) and "generic" (T, S) tensors, nan-means over classes / thresholds as numpy arrays."""
    device = confs.device if device is None else device
    with torch.no_grad():
        ap = _average_precision_all_thresholds(confs.to(device), is_positive.to(device), size_categories.to(device),
                                               num_annos.to(device)).cpu()
    with warnings.catch_warnings():  # all-NaN slices (a size bucket without annotations) are expected
        warnings.simplefilter("ignore", RuntimeWarning)
        return {
            "all": ap[:, 0:-1, :],
            "classavg": np.nanmean(ap[:, 0:-1, :], axis=1),
            "threshavg": np.nanmean(ap[:, 0:-1, :], axis=0),
            "classavg threshavg": np.nanmean(ap[:, 0:-1, :], axis=(0, 1)),
            "generic": ap[:, -1, :],
            "generic threshavg": np.nanmean(ap[:, -1, :], axis=0),
        }


class ApAccumulator:
    """Collects the per-step `od_map_stuffs` of an epoch (trainer.py:155,205-211) and aggregates them at its end.

        acc = ApAccumulator(distributed=world_size > 1)
        for batch in loader:
            ..., od_map_stuffs = model(batch)      # or the static outputs of a replayed CUDA graph
            acc.add(od_map_stuffs)                 # no host sync
        ap = acc.aggregate()                       # the only sync; same dict on every rank, as upstream
    """

    def __init__(self, distributed=False, slab_floats=1 << 22):
        self.distributed = bool(distributed) and dist.is_available() and dist.is_initialized()
        self._slab_floats = int(slab_floats)
        self._slabs, self._used = [], 0
        self._items = []    # (host view (world, numel), event or None, shapes, stat keys)
        self._stream = None
        self.last_stats_device = None
        self.d2h_bytes_last = 0

    def __len__(self):
        return len(self._items)

    def _host_view(self, numel, pinned):
        if not self._slabs or self._used + numel > self._slabs[-1].numel():
            self._slabs.append(torch.empty(max(self._slab_floats, numel), dtype=torch.float32, pin_memory=pinned))
            self._used = 0
        view = self._slabs[-1][self._used:self._used + numel]
        self._used += numel
        return view

    def add(self, stuffs, stats=None):
        """Queue one step's `od_map_stuffs`.  stats: the step's dict of logged scalars (st_detr.py:181-187) — they ride
        in the same packed buffer, so the N > 1 path still costs ONE collective per step (replacing the reference's
        all-reduce of `reduce_distrib_loss`, utils/distributed.py:63-75, plus its four all-gathers, :78-89) and one
        device->host copy.  The mean over ranks is available without a host sync as `last_stats_device` and, once the
        step's copy has landed, from `stats_history()`."""
        confs, is_positive, size_categories, num_annos = stuffs
        shapes = (tuple(confs.shape), tuple(is_positive.shape), tuple(size_categories.shape), tuple(num_annos.shape))
        keys = tuple(sorted(stats)) if stats else ()  # sorted: consistent across processes (distributed.py:69)
        with torch.no_grad():
            parts = [t.detach().reshape(-1).to(torch.float32) for t in (confs, is_positive, size_categories, num_annos)]
            if keys:
                parts.append(torch.stack([stats[k].detach().reshape(()).to(torch.float32) for k in keys]))
            packed = torch.cat(parts)
            if self.distributed:  # one collective for everything; every rank keeps every rank's objects
                gathered = [torch.empty_like(packed) for _ in range(dist.get_world_size())]
                dist.all_gather(gathered, packed)
                packed = torch.stack(gathered)
            else:
                packed = packed[None]
            if keys:
                self.last_stats_device = dict(zip(keys, packed[:, -len(keys):].mean(dim=0)))
            host = self._host_view(packed.numel(), pinned=packed.is_cuda).view(packed.shape)
            event = None
            if packed.is_cuda:
                if self._stream is None:
                    self._stream = torch.cuda.Stream(device=packed.device)
                self._stream.wait_stream(torch.cuda.current_stream(packed.device))
                with torch.cuda.stream(self._stream):
                    host.copy_(packed, non_blocking=True)
                    event = torch.cuda.Event()
                    event.record(self._stream)
                packed.record_stream(self._stream)
            else:
                host.copy_(packed)
        self._items.append((host, event, shapes, keys))
        self.d2h_bytes_last = 4 * packed.numel()

    def stats_history(self):
        """Per step, the logged scalars averaged over ranks (what trainer.py:197-199 feeds its meters), as floats.
        Synchronises on the steps' copy events only."""
        out = []
        for host, event, shapes, keys in self._items:
            if not keys:
                out.append({})
                continue
            if event is not None:
                event.synchronize()
            out.append(dict(zip(keys, host[:, -len(keys):].mean(dim=0).tolist())))
        return out

    def tensors(self):
        """The four epoch tensors in the trainer's order (trainer.py:231-236): objects concatenated step-major with the
        ranks in order inside a step; num_annos stacked to (C, S, steps*world)."""
        if not self._items:
            raise ValueError("ApAccumulator.tensors() called before any add()")
        out = [[], [], [], []]
        for host, event, shapes, _keys in self._items:
            if event is not None:
                event.synchronize()
            for rank_row in host:
                o = 0
                for i, shp in enumerate(shapes):
                    n = int(np.prod(shp))
                    out[i].append(rank_row[o:o + n].view(shp))
                    o += n
        confs = torch.cat(out[0], dim=2)
        is_positive = torch.cat(out[1], dim=2) != 0
        size_categories = torch.cat(out[2], dim=2) != 0
        num_annos = torch.stack(out[3], dim=2).to(torch.int64)
        return confs, is_positive, size_categories, num_annos

    def aggregate(self, device=None):
        return aggregate_mean_average_precision(*self.tensors(), device=device)

    def reset(self):
        """Start the next epoch; the pinned slabs are kept."""
        for _, event, _, _ in self._items:
            if event is not None:
                event.synchronize()
        self._items = []
        self._slabs, self._used = self._slabs[:1], 0
