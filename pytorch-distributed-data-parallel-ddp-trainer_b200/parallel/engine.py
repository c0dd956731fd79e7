"""Data-parallel step engines: what ``DDP(model)`` + ``loss.backward()`` + ``opt.step()`` are in the
reference (reference train_ddp.py:34,196-200), re-designed around a flat arena.

``CpuDataParallel``   gloo path (the reference's CPU fallback): torch ops, bucketed
                      ``dist.all_reduce`` on slices of the gradient arena, in-place SGD.
``B200DataParallel``  B200 path: static kernel sequence (models/simple_cnn.py), gradient buckets
                      all-reduced by the fused all-reduce+SGD kernel over the symmetric-memory
                      communicator, bucket 0 on a side stream so it overlaps the conv backward,
                      the whole step captured once per (batch size, parity) in a CUDA graph.

Both: parameters are broadcast from rank 0 at construction (what DDP's constructor does,
SURVEY.md X13) so all ranks start identical without any seed being set.
"""
from __future__ import annotations

import contextlib

import torch
import torch.distributed as dist

from ..models.simple_cnn import IMG_H, IMG_W, SimpleCNN, SimpleCNNB200
from ..utils.checkpoint import sgd_state_dict
from ..utils.trace import nvtx_range


def _align(n: int, a: int) -> int:
    return (n + a - 1) // a * a


class _EngineBase:
    lr: float
    model: object

    def state_dict(self):
        return self.model.state_dict()

    def optimizer_state_dict(self) -> dict:
        return sgd_state_dict(self.lr, len(self.model.arena.specs))

    def set_lr(self, lr: float) -> None:
        self.lr = float(lr)

    def verify_replicas(self, rank: int, world_size: int, **settings) -> None:
        """Every rank must hold the same parameter list and run the same kernel/collective sequence
        (what DDP's constructor checks with _verify_param_shape_across_processes; here a
        mismatch would not be an error message later but a dead-peer timeout).  One small object
        all-gather at construction."""
        if world_size <= 1:
            return
        mine = (tuple((s.name, tuple(s.shape)) for s in self.model.arena.specs), tuple(sorted(settings.items())))
        everyone = [None] * world_size
        dist.all_gather_object(everyone, mine)
        for r, other in enumerate(everyone):
            if other != mine:
                what = "parameter names/shapes" if other[0] != mine[0] else f"engine settings {dict(other[1])} vs {dict(mine[1])}"
                raise RuntimeError(f"rank {rank}: replica mismatch with rank {r}: {what}")


class CpuDataParallel(_EngineBase):
    def __init__(self, rank: int, world_size: int, lr: float = 0.01, device="cpu"):
        self.rank, self.world_size, self.lr = rank, world_size, float(lr)
        self.device = torch.device(device)
        self.model = SimpleCNN(device=self.device)
        self.buckets = self.model.buckets()
        self.verify_replicas(rank, world_size, engine="cpu")
        self.sync_params()

    @property
    def flat_params(self) -> torch.Tensor:
        return self.model.arena.params

    def sync_params(self, src: int = 0) -> None:
        if self.world_size > 1:
            dist.broadcast(self.model.arena.params.detach(), src=src)

    def step(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        a = self.model.arena
        if self.world_size > 1:
            # each bucket's all-reduce starts the moment its gradients exist and runs on gloo's
            # threads under the rest of the backward pass
            works = []

            def launch(i: int) -> None:
                b = self.buckets[i]
                works.append(dist.all_reduce(a.grads[b.start:b.start + b.numel], async_op=True))

            loss = self.model.forward_backward_staged(x, y, launch)
            for w in works:
                w.wait()
        else:
            loss = self.model.forward_backward(x, y)
        with torch.no_grad():
            a.params.add_(a.grads, alpha=-self.lr / self.world_size)
        return loss

    def read_loss(self, loss) -> float:
        return float(loss)

    def close(self, sync_peers: bool = True) -> None:
        pass


class B200DataParallel(_EngineBase):
    def __init__(self, rank: int, world_size: int, device, batch_cap: int, lr: float = 0.01,
                 use_graph: bool = True, algo: int = 0, timeout_ms: int | None = None,
                 tail_update: str | None = None, multicast: bool | None = None,
                 precision: str | None = None):
        """``precision``: "bf16x3" (default — the reference's numerics or better: 4-byte storage, split
        hi+lo tensor-core operands >= TF32 for the convolutions, fp32 linear layer) or "bf16" (fast
        mode), see SimpleCNNB200.  ``tail_update`` selects how the last (conv1) bucket of the step is updated:
        "push" (default) / "pull" = split-K reduce + all-reduce + SGD in one kernel with the push or
        pull exchange, "split" = separate reduce kernel followed by the generic bucket update.
        Environment override: DDPB_TAIL_UPDATE.  ``multicast`` (default: from four GPUs up; DDPB_NVLS=0/1 forces it) asks for the
        VMM + multicast communicator; with it, buckets above 256 KB on more than two GPUs are
        reduced inside the NVSwitch (multimem.ld_reduce / multimem.st)."""
        from .comm import SymmetricComm
        self.rank, self.world_size, self.lr = rank, world_size, float(lr)
        self.device = torch.device(device)
        torch.cuda.set_device(self.device)
        self.batch_cap = int(batch_cap)
        self.use_graph = use_graph
        self.algo = algo
        import os
        self.tail_update = tail_update or os.environ.get("DDPB_TAIL_UPDATE", "push")
        # timing diagnostics only (the step is then WRONG): skip the update of the named buckets to see
        # which stream is the critical path, e.g. DDPB_ABLATE=conv2 or DDPB_ABLATE=fl,pack
        self._ablate = set(filter(None, os.environ.get("DDPB_ABLATE", "").split(",")))
        # "event" (default): the net.2 bucket update waits for the conv2 data gradient, the last reader of
        # net.2.weight (bf16 mode) / of its tile images (bf16x3 mode) in the step.  "legacy" (diagnostic
        # only): the update follows the weight gradient alone and can overtake the data gradient — a real,
        # timing-dependent race; tests/test_model_gpu.py::test_net2_update_waits_for_the_data_gradient
        # provokes it with DDPB_TEST_DELAY_DGRAD (spin cycles inserted before the data gradient).
        self._test_delay_dgrad = int(os.environ.get("DDPB_TEST_DELAY_DGRAD", "0"))
        self.net2_order = os.environ.get("DDPB_NET2_ORDER", "event")
        if self.net2_order not in ("event", "legacy"):
            raise ValueError(f"DDPB_NET2_ORDER must be event or legacy, got {self.net2_order!r}")
        if self.tail_update not in ("push", "pull", "split"):
            raise ValueError(f"tail_update must be push, pull or split, got {self.tail_update!r}")
        # programmatic dependent launch pays off for latency-bound (small-batch) steps only
        from ..ops import native as _nat
        if "DDPB_PDL" not in os.environ:
            _nat.lib().ddpb_set_pdl(1 if self.batch_cap <= 512 else 0)

        from ..models.arena import ParamArena
        from ..models.simple_cnn import SPECS
        n = ParamArena(SPECS, device="meta").padded_numel
        self._n = n
        region = _align(n * 4, 256)
        self.param_off, self.grad_off = 0, [region, 2 * region]
        self.comm = SymmetricComm(rank, world_size, 3 * region, timeout_ms=timeout_ms,
                                  multicast=multicast)
        params = self.comm.tensor(self.param_off, n)
        grads = [self.comm.tensor(off, n) for off in self.grad_off]
        self.model = SimpleCNNB200(batch_cap, self.device, params=params, grads=grads[0],
                                   alt_grads=grads[1], precision=precision)
        self.precision = self.model.precision
        self.model.lr_dev.fill_(self.lr)
        # gradient buckets = contiguous arena slices (whole float4; the arena tail is zero padding),
        # one per point of the backward pass at which a group of gradients becomes complete
        a = self.model.arena
        self.bucket_fl = self._slice(a, ("fl.weight", "fl.bias"))
        self.bucket_conv2 = self._slice(a, ("net.2.weight", "net.2.bias"))
        self.bucket_conv1 = self._slice(a, ("net.0.weight", "net.0.bias"))
        self.buckets = [self.bucket_fl, self.bucket_conv2, self.bucket_conv1]
        assert all(s % 4 == 0 and s + m <= n for s, m in self.buckets)
        self.verify_replicas(rank, world_size, engine="b200", tail_update=self.tail_update, algo=self.algo,
                             multicast=self.comm.multicast, net2_order=self.net2_order,
                             precision=self.precision,
                             ablate=",".join(sorted(self._ablate)))

        # Stream priorities (lower number = its blocks are dispatched first): the two bucket streams run
        # above the main stream, so a bucket's exchange never queues behind the long convolution backward
        # for SM slots (2 GPUs, batch 64: 0.1130 -> 0.1088 ms/step; raising the main stream instead:
        # 0.1184).  DDPB_STREAM_PRIO="main,side,side2" overrides.
        prio = [int(v) for v in os.environ.get("DDPB_STREAM_PRIO", "0,-1,-1").split(",")]
        self.main = torch.cuda.Stream(self.device, priority=prio[0])
        self.side = torch.cuda.Stream(self.device, priority=prio[1])
        self.side2 = torch.cuda.Stream(self.device, priority=prio[2])
        self.copy = torch.cuda.Stream(self.device)
        self._ev_fork = torch.cuda.Event()
        self._ev_loss_done = torch.cuda.Event()
        self._ev_join = torch.cuda.Event()
        self._ev_join2 = torch.cuda.Event()
        self._ev_dgrad = torch.cuda.Event()
        self._ev_ready = [torch.cuda.Event(), torch.cuda.Event()]
        self._ev_free = [torch.cuda.Event(), torch.cuda.Event()]
        self.x_in = [torch.zeros(batch_cap, 1, IMG_H, IMG_W, device=self.device) for _ in range(2)]
        self.y_in = [torch.zeros(batch_cap, dtype=torch.int64, device=self.device) for _ in range(2)]
        self.loss_host = torch.zeros(2, dtype=torch.float32).pin_memory()
        self._ev_loss = [torch.cuda.Event(), torch.cuda.Event()]
        self._graphs: dict = {}
        self._seen: dict = {}
        self.step_count = 0
        self.sync_params()

    # ------------------------------------------------------------------ parameters
    @staticmethod
    def _slice(arena, names):
        start, numel = arena.slice_of(names)
        return start, _align(numel, 4)

    @property
    def flat_params(self) -> torch.Tensor:
        return self.model.arena.params

    def set_lr(self, lr: float) -> None:
        self.lr = float(lr)
        self.model.lr_dev.fill_(self.lr)

    def sync_params(self, src: int = 0) -> None:
        """Initial (and post-resume) parameter broadcast from ``src`` over the peer-mapped arena."""
        torch.cuda.synchronize(self.device)
        if self.world_size > 1:
            self.comm.broadcast(self.param_off, self._n * 4, root=src, slot=7, stream=self.main)
        self.model.pack_weights(self.main)
        self.main.synchronize()

    # ------------------------------------------------------------------ the step
    def launches_per_step(self, batch: int) -> int:
        # the model counts the conv2 split-K reduce as its own kernel and no conv1 reduce; on top of it:
        # three bucket updates + the weight re-pack.  "push" folds the conv2 reduce into its update
        # kernel as well (-1), "split" runs the conv1 reduce separately (+1).
        extra = {"push": -1, "pull": 0, "split": 1}[self.tail_update]
        # bf16x3 mode has no fl.weight re-pack kernel; the net.2 tile re-pack is inside the fused update
        # ("push") or one more small kernel otherwise
        repack = (0 if self.tail_update == "push" else 1) if self.model.x3 else 1
        return self.model.launches_per_step(batch) + 3 + repack + extra

    @contextlib.contextmanager
    def _plain_launch(self):
        """Kernels launched inside are ordinary stream-ordered launches (no programmatic dependent
        launch attribute), whatever the engine-wide setting is."""
        lib = self.model.nat.lib()
        prev = lib.ddpb_get_pdl()
        lib.ddpb_set_pdl(0)
        try:
            yield
        finally:
            lib.ddpb_set_pdl(prev)

    def _update(self, bucket, parity: int, slot: int, stream) -> None:
        """Fused all-reduce + 1/W + SGD of one gradient bucket (plain SGD kernel when W == 1)."""
        m = self.model
        start, n = bucket
        if self.world_size > 1:
            goff = self.grad_off[parity]
            self.comm.allreduce_sgd(goff + start * 4, self.param_off + start * 4, n, m.lr_dev,
                                    self.algo, slot=slot, stream=stream)
        else:
            m.nat.sgd(m.arena.params[start:start + n], m.grad_sets[parity]["flat"][start:start + n],
                      m.lr_dev, stream)

    def _enqueue(self, b: int, parity: int) -> None:
        """Enqueue one full training step; capturable.  Stream plan (each update is launched as
        soon as its gradients are complete, so only the tiny conv1 bucket is left at the end):

            main : fwd+loss, linear bwd | conv2 dgrad -> conv1 wgrad -> reduce+update {net.0.*} (one kernel)
            side : (after linear bwd)    update {fl.*} -> re-pack fl.weight tiles
            side2: (after linear bwd)    conv2 wgrad -> [waits for dgrad] -> reduce+update {net.2.*}

        A bucket's update rewrites its parameters in place, so it must come after the last kernel of the
        step that reads them: fl.* is last read by the linear backward (before the fork), net.0.* by the
        conv1 forward, net.2.weight by the conv2 data gradient on the main stream — hence the event.
        """
        m, main = self.model, self.main
        x, y = self.x_in[parity][:b], self.y_in[parity][:b]
        with nvtx_range("fwd+loss"):
            m.forward_loss(x, y, b, main, parity, pack=False)
        with nvtx_range("bwd linear"):
            if m.x3:
                # the dW half of the linear backward needs only act2 and dlogits: its stream forks right
                # after the loss, so the fl bucket's chain (dW half -> reduce -> exchange+SGD) starts while
                # the dX half still runs here
                self._ev_loss_done.record(main)
                self.side.wait_event(self._ev_loss_done)
            m.backward_head(b, main, parity, defer_reduce=True)
        self._ev_fork.record(main)
        if not m.x3:
            self.side.wait_event(self._ev_fork)
        self.side2.wait_event(self._ev_fork)
        with nvtx_range("bucket fl: all-reduce+sgd, re-pack"):
            m.backward_head_reduce(self.side, parity)   # bf16x3: dW half + its split-partial sum
            if m.x3:
                # fl.weight may only be rewritten once its last reader, the dX half on the main stream,
                # is done; the update then has two predecessors -> ordinary (non-programmatic) launch
                self.side.wait_event(self._ev_fork)
            if "fl" not in self._ablate:
                with (self._plain_launch() if m.x3 else contextlib.nullcontext()):
                    self._update(self.bucket_fl, parity, 0, self.side)
            if "pack" not in self._ablate and not m.x3:
                m.pack_weights(self.side)  # new fl.weight -> tensor-core tiles for the next step
        self._ev_join.record(self.side)
        # side2: conv2 weight gradient (shares the SMs with the data gradient below: 112 KB + 112 KB of
        # shared memory, 384 + 64 TMEM columns)
        fused2 = self.tail_update == "push" and "conv2" not in self._ablate
        with nvtx_range("bwd conv2 wgrad"):
            if fused2:
                nparts2 = m.backward_conv2_wgrad_partials(b, self.side2)
            else:
                m.backward_conv2_wgrad(b, self.side2, parity)
        # main: conv2 data gradient, the last reader of net.2.weight in this step
        with nvtx_range("bwd conv2 dgrad"):
            if self._test_delay_dgrad:
                torch.cuda._sleep(self._test_delay_dgrad)  # test hook: hold the data gradient back
            m.backward_dgrad(b, main)
        # side2: net.2.* may only be rewritten once the data gradient is done with it.  The update is
        # then the one kernel of the step with two predecessors (wgrad in stream order, dgrad through
        # the event), so it is launched with ordinary full dependencies, not as a programmatic dependent.
        ordered = self.net2_order == "event"
        if ordered:
            self._ev_dgrad.record(main)
            self.side2.wait_event(self._ev_dgrad)
        with nvtx_range("bucket net.2"), (self._plain_launch() if ordered else contextlib.nullcontext()):
            if fused2:
                # same one-kernel split-K reduce + exchange + SGD as the tail bucket below
                from .comm import ALGO_PUSH, CONV3X3_WGRAD_IMAGE, LAYOUT_CONV3X3_WGRAD
                s2, _n2 = self.bucket_conv2
                self.comm.reduce_allreduce_sgd(
                    m.scr_w2, nparts2, CONV3X3_WGRAD_IMAGE, self.grad_off[parity] + s2 * 4,
                    self.param_off + s2 * 4, m.lr_dev, algo=ALGO_PUSH, layout=LAYOUT_CONV3X3_WGRAD,
                    slot=2, stream=self.side2, pack=m.w2_packed,  # bf16x3: tiles rewritten in the same kernel
                    bias_parts=(m.db2_part, *m.db2_extent, IMG_H * IMG_W) if m.x3 else None)
            elif "conv2" not in self._ablate:
                self._update(self.bucket_conv2, parity, 2, self.side2)
                if m.x3:
                    m.pack_weights(self.side2)  # new net.2.weight -> split tiles for the next step
        self._ev_join2.record(self.side2)
        with nvtx_range("bwd conv1 wgrad + bucket net.0"):
            if self.tail_update == "split":
                m.backward_conv1_wgrad(x, b, main, parity)
                if "conv1" not in self._ablate:
                    self._update(self.bucket_conv1, parity, 1, main)
            elif "conv1" in self._ablate:
                m.backward_conv1_wgrad_partials(x, b, main)
            else:
                from .comm import ALGO_ONESHOT, ALGO_PUSH
                nparts = m.backward_conv1_wgrad_partials(x, b, main)
                s1, n1 = self.bucket_conv1
                # split-K reduce of the conv1 partials + all-reduce + 1/W + SGD in ONE kernel: this
                # bucket is the tail of the step's critical path
                self.comm.reduce_allreduce_sgd(
                    m.scr_w1, nparts, n1, self.grad_off[parity] + s1 * 4, self.param_off + s1 * 4,
                    m.lr_dev, algo=ALGO_PUSH if self.tail_update == "push" else ALGO_ONESHOT, slot=1,
                    stream=main)
        main.wait_event(self._ev_join2)
        main.wait_event(self._ev_join)

    def acquire_inputs(self):
        """Device buffers the next step() will read.  A loader may fill them in place on the
        current stream: that stream is first made to wait for the step that last read them."""
        p = self.step_count & 1
        torch.cuda.current_stream(self.device).wait_event(self._ev_free[p])
        return self.x_in[p], self.y_in[p]

    def state_dict(self):
        self.synchronize()
        return self.model.state_dict()

    def step(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        """One data-parallel training step.  ``x`` float32 [b,1,28,28], ``y`` int64 [b]; CPU
        (ideally pinned) or CUDA tensors.  Returns the device loss scalar (valid on self.main;
        use read_loss() for a synchronised Python float)."""
        b = int(x.shape[0])
        parity = self.step_count & 1
        xin, yin = self.x_in[parity], self.y_in[parity]
        if x.data_ptr() != xin.data_ptr():
            if x.is_cuda:
                cur = torch.cuda.current_stream(self.device)
                self.main.wait_stream(cur)
                with torch.cuda.stream(self.main):
                    xin[:b].copy_(x, non_blocking=True)
                    yin[:b].copy_(y, non_blocking=True)
            else:
                with torch.cuda.stream(self.copy):
                    self.copy.wait_event(self._ev_free[parity])
                    xin[:b].copy_(x, non_blocking=True)
                    yin[:b].copy_(y, non_blocking=True)
                    self._ev_ready[parity].record(self.copy)
                self.main.wait_event(self._ev_ready[parity])
        else:
            self.main.wait_stream(torch.cuda.current_stream(self.device))
        key = (b, parity)
        seen = self._seen.get(key, 0)
        self._seen[key] = seen + 1
        if self.use_graph and seen >= 1:
            g = self._graphs.get(key)
            if g is None:
                g = torch.cuda.CUDAGraph()
                self.main.synchronize()
                with torch.cuda.graph(g, stream=self.main):
                    self._enqueue(b, parity)
                self._graphs[key] = g
            with torch.cuda.stream(self.main), nvtx_range("step graph replay"):
                g.replay()
        else:
            with torch.cuda.stream(self.main):
                self._enqueue(b, parity)
        with torch.cuda.stream(self.main):
            self.loss_host[parity:parity + 1].copy_(self.model.loss, non_blocking=True)
            self._ev_loss[parity].record(self.main)
            self._ev_free[parity].record(self.main)
        self.step_count += 1
        return self.model.loss

    def read_loss(self, loss=None) -> float:
        """Loss of the most recent step as a Python float (waits for that step only)."""
        parity = (self.step_count - 1) & 1
        self._ev_loss[parity].synchronize()
        return float(self.loss_host[parity])

    def synchronize(self) -> None:
        self.main.synchronize()
        self.side.synchronize()
        self.side2.synchronize()
        self.copy.synchronize()

    def close(self, sync_peers: bool = True) -> None:
        """Release the graphs and the peer-mapped buffer.  Peers read this rank's buffer from inside
        their kernels, so it must not be freed while any rank still has a step in flight: with
        ``sync_peers`` (default) all ranks meet at a host barrier after draining their own streams.
        Pass False only on an error path where the peers may never arrive."""
        self.synchronize()
        if sync_peers and self.world_size > 1 and dist.is_available() and dist.is_initialized():
            dist.barrier()
        self._graphs.clear()
        self.comm.close()


def make_engine(rank: int, world_size: int, device, batch_cap: int, lr: float = 0.01, **kw):
    """CUDA device -> B200DataParallel, CPU -> CpuDataParallel (the reference's backend switch,
    reference utils.py:6)."""
    device = torch.device(device)
    if device.type == "cuda":
        return B200DataParallel(rank, world_size, device, batch_cap, lr=lr, **kw)
    return CpuDataParallel(rank, world_size, lr=lr, device=device)
