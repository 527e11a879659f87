"""Training-step drivers for the drop-in model: the reference Trainer's inner step (future_od/trainer.py:171-189:
zero_grad, forward, backward, clip_grad_norm_(0.1), optimizer.step) either eagerly or as ONE CUDA graph.

The whole step is capturable because nothing in it touches the host: the matcher / loss run on the device
(no `.cpu()`, no scipy, no `.item()`), shapes are static, and AdamW runs fused with capturable state.
"""
import torch


def setup_optimizer(model, detr_args, capturable=False):
    """runs/_helper.py:84-107: AdamW with the backbone parameters in their own group at lr_backbone."""
    named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
    groups = [{"params": [p for n, p in named if "backbone" not in n]},
              {"params": [p for n, p in named if "backbone" in n], "lr": detr_args.lr_backbone}]
    return torch.optim.AdamW(groups, lr=detr_args.lr, weight_decay=detr_args.weight_decay, fused=True,
                             capturable=capturable)


class TrainStep:
    """Eager step.  `model` may be DDP-wrapped; `params` are the trainable parameters that get clipped.

    With `reducer` (fod_b200.ddp.BucketedGradReducer over the UNWRAPPED model's parameters) the gradients live in flat
    fp32 buckets that are averaged over ranks during backward and clipped bucket-wise; this form has static structure
    and is what GraphedTrainStep captures for N > 1 GPUs."""

    def __init__(self, model, optimizer, max_norm=0.1, distributed=False, reducer=None):
        """optimizer: a torch optimizer, or a fod_b200.optim.FlatTrainState (flat arenas: gradient average, clipping,
        AdamW, bf16 operand refresh and zero_grad fused; `reducer` is then its own)."""
        self.model, self.opt, self.max_norm, self.distributed = model, optimizer, max_norm, distributed
        self.params = [p for p in model.parameters() if p.requires_grad]
        self.reducer = reducer
        self.flat = optimizer if hasattr(optimizer, "finish_backward") else None
        self.stats = self.od_map_stuffs = None  # the step's logged scalars / AP bookkeeping (trainer.py:191-211)

    def __call__(self, data):
        if self.flat is not None:
            self.flat.begin_step()
            _, _, loss, self.stats, self.od_map_stuffs = self.model(data=data, distributed=self.distributed)
            loss.backward()
            self.flat.finish_backward()
            self.flat.step(self.max_norm)
            return loss
        if self.reducer is not None:
            self.reducer.zero_grad()
        else:
            self.opt.zero_grad(set_to_none=True)
        _, _, loss, self.stats, self.od_map_stuffs = self.model(data=data, distributed=self.distributed)
        loss.backward()
        if self.reducer is not None:
            self.reducer.finish()
            self.reducer.clip_grad_norm_(self.max_norm)
        else:
            torch.nn.utils.clip_grad_norm_(self.params, self.max_norm)
        self.opt.step()
        return loss


class GraphedTrainStep:
    """The same step captured into one CUDA graph.  Inputs are copied into static device buffers, the graph is
    replayed, and the static loss tensor is returned (read it with .item() if needed).  With N > 1 ranks pass a
    `reducer`: the bucketed NCCL all-reduces and the num_boxes exchange become nodes of the graph (side-stream
    branches that overlap backward), so a multi-GPU step is still ONE cudaGraphLaunch per rank.

    The optimizer must have been built with capturable=True.  `warmup` eager steps run first on a side stream, as
    CUDA-graph capture requires (allocator warm-up, cuDNN autotuning, lazily created kernel attributes).  They are
    steps on the EXAMPLE batch that the reference's Trainer never takes, so the trainable state (parameters, both
    AdamW moments, the step counter) is snapshotted before them and restored afterwards (`keep_warmup_updates=True`
    keeps them): the first replay is the first optimizer step, with bias correction at t = 1."""

    def __init__(self, model, optimizer, example, max_norm=0.1, warmup=3, distributed=False, reducer=None,
                 keep_warmup_updates=False):
        self.eager = TrainStep(model, optimizer, max_norm, distributed=distributed, reducer=reducer)
        self.static = {k: (v.clone() if torch.is_tensor(v) else v) for k, v in example.items()}
        snap = None if keep_warmup_updates else self._snapshot(model, optimizer)
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(warmup):
                self.eager(self.static)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        if snap is not None:
            self._restore(model, optimizer, snap)
            del snap
        self.graph = torch.cuda.CUDAGraph()
        if reducer is None and hasattr(optimizer, "zero_grad"):
            optimizer.zero_grad(set_to_none=True)
        # thread_local: the NCCL watchdog thread may query events while this thread captures
        with torch.cuda.graph(self.graph, capture_error_mode="thread_local" if distributed else "global"):
            self.loss = self.eager(self.static)
        torch.cuda.synchronize()
        self._lr_version = getattr(self.eager.flat, "lr_version", 0)

    @staticmethod
    def _snapshot(model, optimizer):
        with torch.no_grad():
            if hasattr(optimizer, "finish_backward"):  # FlatTrainState: four arenas
                return ("flat", optimizer.params.clone(), optimizer.exp_avg.clone(), optimizer.exp_avg_sq.clone(),
                        optimizer.state.clone())
            import copy
            return ("torch", [p.detach().clone() for p in model.parameters() if p.requires_grad],
                    copy.deepcopy(optimizer.state_dict()))

    @staticmethod
    def _restore(model, optimizer, snap):
        with torch.no_grad():
            if snap[0] == "flat":
                optimizer.params.copy_(snap[1])
                optimizer.exp_avg.copy_(snap[2])
                optimizer.exp_avg_sq.copy_(snap[3])
                optimizer.state.copy_(snap[4])
                optimizer.grads.zero_()
                optimizer.refresh_shadow()
            else:
                for p, q in zip([p for p in model.parameters() if p.requires_grad], snap[1]):
                    p.copy_(q)
                optimizer.load_state_dict(snap[2])
        torch.cuda.synchronize()

    def __call__(self, data=None):
        """data: a batch (host or device tensors) copied into the static inputs now, or None to consume the oldest
        batch queued with prefetch()."""
        cur = torch.cuda.current_stream()
        if data is None:
            j, ready = self._ready.pop(0)
            cur.wait_event(ready)
            for k, v in self._staging[j].items():
                self.static[k].copy_(v, non_blocking=True)
            self._consumed[j] = torch.cuda.Event()
            self._consumed[j].record(cur)
        else:
            for k, v in data.items():
                if torch.is_tensor(v):
                    self.static[k].copy_(v, non_blocking=True)
        flat = self.eager.flat
        if flat is not None and getattr(flat, "lr_version", 0) != getattr(self, "_lr_version", 0):
            self.recapture()  # FlatTrainState.set_lr() was called: the rates are kernel arguments inside the graph
        self.graph.replay()
        return self.loss

    def recapture(self):
        """Capture the step again (same static buffers, warm allocator): needed after anything that changes kernel
        ARGUMENTS baked into the graph — the learning rates of the fused AdamW kernel (FlatTrainState.set_lr, the
        reference's StepLR, runs/_helper.py:104-106), `max_norm`.  Called automatically for learning-rate changes."""
        torch.cuda.synchronize()
        eager = self.eager
        # The old graph's private pool holds a full step of activations: it is released BEFORE the new capture so the
        # two never coexist (B = 32 at 448x800 / B = 16 at 896x1600 would otherwise transiently need twice the memory).
        # The static outputs (`loss`, `eager.stats`, `eager.od_map_stuffs`) are new tensors afterwards.
        self.graph = self.loss = None
        eager.stats = eager.od_map_stuffs = None
        import gc
        gc.collect()
        graph = torch.cuda.CUDAGraph()
        if eager.reducer is None and hasattr(eager.opt, "zero_grad"):
            eager.opt.zero_grad(set_to_none=True)
        with torch.cuda.graph(graph, capture_error_mode="thread_local" if eager.distributed else "global"):
            loss = eager(self.static)
        torch.cuda.synchronize()
        self.graph, self.loss = graph, loss
        self._lr_version = getattr(eager.flat, "lr_version", 0)

    def read_loss_async(self):
        """Queue a copy of the loss of the step launched LAST into pinned host memory (side stream, after that step) and
        return (host tensor, event): `event.synchronize(); float(host)` reads it without draining the compute stream,
        so the next step can be launched first.  A loop that logs step i's loss after launching step i+1 hides the
        launch latency of the ~1 500-node graph (~1 ms), which a synchronous `float(loss)` per step exposes."""
        if not hasattr(self, "_loss_stream"):
            self._loss_stream = torch.cuda.Stream()
            self._loss_host = [torch.empty((), dtype=torch.float32).pin_memory() for _ in range(4)]
            self._loss_i = 0
        host = self._loss_host[self._loss_i % 4]
        self._loss_i += 1
        self._loss_stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self._loss_stream):
            host.copy_(self.loss.detach().float(), non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._loss_stream)
        return host, ev

    def prefetch(self, host_data):
        """Queue the NEXT batch: pinned host -> device staging buffer on a copy stream, overlapping the step that is
        running.  Two staging slots; a slot is rewritten only after the step that consumed it has copied it out."""
        if not hasattr(self, "_staging"):
            self._copy_stream = torch.cuda.Stream()
            self._staging = [{k: torch.empty_like(v) for k, v in self.static.items() if torch.is_tensor(v)} for _ in range(2)]
            self._consumed, self._ready, self._slot = [None, None], [], 0
        assert len(self._ready) < 2, "at most two batches in flight"
        j = self._slot
        self._slot ^= 1
        with torch.cuda.stream(self._copy_stream):
            if self._consumed[j] is not None:
                self._copy_stream.wait_event(self._consumed[j])
            for k, v in self._staging[j].items():
                v.copy_(host_data[k], non_blocking=True)
            ready = torch.cuda.Event()
            ready.record(self._copy_stream)
        self._ready.append((j, ready))
