"""Symmetric-memory communicator (Python side).

Owns this rank's peer-mapped buffer (native: csrc/comm.cu) and exposes slices of it as torch
tensors, plus the in-kernel-flag collectives that replace what the reference gets from
ProcessGroupNCCL + the c10d Reducer (reference utils.py:7-11, train_ddp.py:34,199;
SURVEY.md §2.5 X13, X16, §5.9).  torch.distributed is used once, as plumbing, to exchange the
64-byte CUDA IPC handles.

Two ways of sharing the buffer:

* CUDA IPC (default): ``cudaMalloc`` + ``cudaIpcGetMemHandle``; peers open the handle.
* VMM + multicast (``multicast=True`` or ``DDPB_NVLS=1``): the buffer is a ``cuMemCreate``
  allocation exported as a POSIX file descriptor; descriptors travel between the rank processes
  over AF_UNIX sockets (SCM_RIGHTS), every rank maps every peer, and all the allocations are bound
  to ONE multicast object so that ``multimem.ld_reduce`` / ``multimem.st`` reduce and broadcast
  inside the NVSwitch (NVLS, ``ALGO_NVLS``).  Every set-up stage is agreed on by all ranks; if any
  rank fails a stage, all of them fall back to the IPC communicator.

All ranks must issue the same sequence of collective calls per flag slot.
"""
from __future__ import annotations

import ctypes
import os
import socket
import threading
import uuid
import warnings
from ctypes import c_int, c_longlong, c_size_t, c_uint, c_void_p, POINTER

import torch
import torch.distributed as dist

from ..ops import native as nat

P, I, LL, SZ = c_void_p, c_int, c_longlong, c_size_t

for _name, _res, _args in (
    ("ddpb_comm_header_bytes", SZ, []),
    ("ddpb_comm_handle_bytes", I, []),
    ("ddpb_comm_max_world", I, []),
    ("ddpb_comm_num_slots", I, []),
    ("ddpb_comm_create", I, [I, I, SZ, POINTER(c_void_p)]),
    ("ddpb_comm_get_handle", I, [P, P]),
    ("ddpb_comm_connect", I, [P, P]),
    ("ddpb_comm_local_ptr", c_void_p, [P]),
    ("ddpb_comm_peer_ptr", c_void_p, [P, I]),
    ("ddpb_comm_payload_bytes", SZ, [P]),
    ("ddpb_comm_set_timeout_ms", None, [P, LL]),
    ("ddpb_comm_status", I, [P, POINTER(c_uint)]),
    ("ddpb_comm_destroy", I, [P]),
    ("ddpb_multicast_supported", I, []),
    ("ddpb_comm_create_vmm", I, [I, I, SZ, POINTER(c_void_p)]),
    ("ddpb_comm_export_fd", I, [P, POINTER(c_int)]),
    ("ddpb_comm_import_peer_fd", I, [P, I, I]),
    ("ddpb_comm_mc_create", I, [P, POINTER(c_int)]),
    ("ddpb_comm_mc_import", I, [P, I]),
    ("ddpb_comm_mc_add_device", I, [P]),
    ("ddpb_comm_mc_bind", I, [P]),
    ("ddpb_comm_set_connected", I, [P]),
    ("ddpb_comm_has_multicast", I, [P]),
    ("ddpb_comm_multicast_ptr", c_void_p, [P]),
    ("ddpb_allreduce_sgd", I, [P, SZ, SZ, LL, P, I, I, P]),
    ("ddpb_reduce_allreduce_sgd", I, [P, P, I, LL, SZ, SZ, P, I, I, I, P]),
    ("ddpb_reduce_allreduce_sgd_pack", I, [P, P, I, LL, SZ, SZ, P, I, I, I, P, P, I, I, I, P]),
    ("ddpb_allreduce", I, [P, SZ, SZ, LL, I, I, I, P]),
    ("ddpb_broadcast", I, [P, SZ, LL, I, I, P]),
):
    nat.declare(_name, _res, _args)

ALGO_AUTO, ALGO_ONESHOT, ALGO_TWOSHOT, ALGO_PUSH, ALGO_NVLS = 0, 1, 2, 3, 4
# one NVSwitch node; the kernels are instantiated for every world size 1..8 (csrc/comm.cu)
SUPPORTED_WORLD_SIZES = (1, 2, 3, 4, 5, 6, 7, 8)
DTYPE_CODES = {torch.float32: 0, torch.bfloat16: 1}
# how partial element i maps into the gradient/parameter slice of reduce_allreduce_sgd
LAYOUT_IDENTITY, LAYOUT_CONV3X3_WGRAD = 0, 1
CONV3X3_WGRAD_IMAGE = 64 * (9 * 32 + 8)  # floats per partial for LAYOUT_CONV3X3_WGRAD


def conv3x3_wgrad_image_dest() -> torch.Tensor:
    """Destination index (into net.2.weight ++ net.2.bias, -1 = unused column) of every element of a
    conv3x3 <32,64> weight-gradient accumulator image; the Python restatement of ll_dest<1> in
    csrc/comm.cu, used by the tests."""
    CI, CO = 32, 64
    n0, n1 = 3 * CI + 8, 3 * CI
    ncols = n0 + 2 * n1
    idx = torch.arange(CO * ncols)
    co, col = idx // ncols, idx % ncols
    dest = torch.full_like(idx, -1)
    r = torch.where(col < n0, 0, torch.where(col < n0 + n1, 1, 2))
    c = torch.where(col < n0, col, torch.where(col < n0 + n1, col - n0, col - n0 - n1))
    tap = (r > 0) | (col < n1)
    sx, ci = c // CI, c % CI
    dest[tap] = ((co * CI + ci) * 9 + r * 3 + sx)[tap]
    bias = (r == 0) & (col == n1)
    dest[bias] = (CO * CI * 9 + co)[bias]
    return dest


class _RawCudaBuffer:
    """Adapter so torch can wrap a raw device pointer without copying."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2, "strides": None,
        }


def exchange_fds(rank: int, world_size: int, fds: list[int], token: str) -> list[list[int]]:
    """All-to-all hand-over of open file descriptors between the rank processes of one node.

    Every rank serves ``fds`` (possibly empty) on an abstract-namespace AF_UNIX socket named after
    ``token`` and its rank, and fetches every peer's list.  Returns ``got[r]`` = descriptors
    received from rank r (new numbers in this process; ``got[rank]`` is empty).  The caller closes
    what it receives.  Needs an initialised torch.distributed group for the two barriers."""
    name = lambda r: f"\0ddpb-{token}-{r}"
    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    srv.bind(name(rank))
    srv.listen(world_size)
    err: list[BaseException] = []

    def serve():
        try:
            for _ in range(world_size - 1):
                conn, _addr = srv.accept()
                with conn:
                    # one data byte carries the count; the descriptors ride along as SCM_RIGHTS
                    socket.send_fds(conn, [bytes([len(fds)])], fds)
        except BaseException as e:  # surfaced by the caller after join()
            err.append(e)

    th = threading.Thread(target=serve, daemon=True)
    th.start()
    dist.barrier()  # every socket is listening
    got: list[list[int]] = [[] for _ in range(world_size)]
    try:
        for r in range(world_size):
            if r == rank:
                continue
            with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as c:
                c.settimeout(60)
                c.connect(name(r))
                msg, rfds, _flags, _addr = socket.recv_fds(c, 1, 16)
                if len(msg) != 1 or msg[0] != len(rfds):
                    for fd in rfds:
                        os.close(fd)
                    raise RuntimeError(f"descriptor hand-over from rank {r} was truncated")
                got[r] = list(rfds)
    finally:
        th.join(timeout=60)
        srv.close()
    if err:
        raise err[0]
    dist.barrier()
    return got


def _all_ok(ok: bool, world_size: int) -> bool:
    """True only if every rank passes True (one tiny collective per set-up stage)."""
    if world_size == 1:
        return ok
    flags = [None] * world_size
    dist.all_gather_object(flags, bool(ok))
    return all(flags)


class SymmetricComm:
    def __init__(self, rank: int, world_size: int, payload_bytes: int, timeout_ms: int | None = None,
                 multicast: bool | None = None):
        if world_size not in SUPPORTED_WORLD_SIZES or not 0 <= rank < world_size:
            raise ValueError(f"the peer-memory communicator is built for world sizes {SUPPORTED_WORLD_SIZES} "
                             f"(one NVSwitch node); got rank {rank} of {world_size}")
        lib = nat.lib()
        self.lib = lib
        self.rank, self.world_size = rank, world_size
        self._h = None
        self.multicast = False
        if multicast is None:
            # Default: the multicast (NVLS) communicator from four GPUs up — measured on 8 B200s at
            # per-rank batch 64: 0.120 ms/step against 0.137 with CUDA IPC, and it wins the all-reduce
            # sweep at every size.  DDPB_NVLS=0/1 forces either.  Where multicast
            # cannot be set up, all ranks agree to fall back to CUDA IPC (RuntimeWarning on rank 0).
            env = os.environ.get("DDPB_NVLS")
            multicast = (env == "1") if env is not None else world_size >= 4
        if multicast and world_size > 1:
            self._h = self._create_vmm(rank, world_size, payload_bytes)
        if self._h is None:
            self._h = self._create_ipc(rank, world_size, payload_bytes)
        if timeout_ms is not None:
            lib.ddpb_comm_set_timeout_ms(self._h, int(timeout_ms))
        self.payload_bytes = int(lib.ddpb_comm_payload_bytes(self._h))
        self._base_ptr = int(lib.ddpb_comm_local_ptr(self._h))
        self._keep = _RawCudaBuffer(self._base_ptr, self.payload_bytes)
        self.buffer = torch.as_tensor(self._keep, device=torch.device("cuda", torch.cuda.current_device()))

    # ------------------------------------------------------------------ set-up
    def _create_ipc(self, rank: int, world_size: int, payload_bytes: int):
        lib = self.lib
        handle = c_void_p()
        nat.check(lib.ddpb_comm_create(rank, world_size, payload_bytes, ctypes.byref(handle)),
                  "comm_create")
        if world_size > 1:
            hb = lib.ddpb_comm_handle_bytes()
            mine = ctypes.create_string_buffer(hb)
            nat.check(lib.ddpb_comm_get_handle(handle, mine), "comm_get_handle")
            gathered = [None] * world_size
            dist.all_gather_object(gathered, bytes(mine.raw))
            blob = b"".join(gathered)
            nat.check(lib.ddpb_comm_connect(handle, blob), "comm_connect (CUDA IPC open)")
            dist.barrier()
        return handle

    def _create_vmm(self, rank: int, world_size: int, payload_bytes: int):
        """VMM allocation + descriptor exchange + multicast bind.  Returns the native handle, or
        None (after releasing everything) when any rank failed any stage."""
        lib = self.lib
        why = ""
        handle = c_void_p()
        rc = lib.ddpb_comm_create_vmm(rank, world_size, payload_bytes, ctypes.byref(handle))
        ok = rc == 0 and lib.ddpb_multicast_supported() == 1
        if rc != 0:
            why = f"create_vmm: {nat.error_string(rc)}"
            handle = None
        elif not ok:
            why = "device does not support multicast objects"

        def give_up(reason: str):
            if handle is not None:
                lib.ddpb_comm_destroy(handle)
            if rank == 0:
                warnings.warn(f"ddp_b200: multicast communicator unavailable ({reason or 'a peer failed'}); "
                              "using the CUDA IPC communicator", RuntimeWarning, stacklevel=3)
            return None

        if not _all_ok(ok, world_size):
            return give_up(why)
        tok = [uuid.uuid4().hex if rank == 0 else None]
        dist.broadcast_object_list(tok, src=0)
        received: list[int] = []
        try:
            # stage 1: every rank maps every peer's allocation
            fd = c_int(-1)
            rc = lib.ddpb_comm_export_fd(handle, ctypes.byref(fd))
            ok, why = rc == 0, (f"export_fd: {nat.error_string(rc)}" if rc else "")
            mine = [fd.value] if ok else []
            got = exchange_fds(rank, world_size, mine, tok[0] + "m")
            for f in mine:
                os.close(f)
            for r, lst in enumerate(got):
                received += lst
                if r == rank:
                    continue
                if len(lst) != 1:
                    ok, why = False, why or f"rank {r} exported no descriptor"
                    continue
                rc = lib.ddpb_comm_import_peer_fd(handle, r, lst[0])
                if rc != 0:
                    ok, why = False, why or f"import_peer_fd({r}): {nat.error_string(rc)}"
            if not _all_ok(ok, world_size):
                return give_up(why)
            # stage 2: rank 0 creates the multicast object, the others import it
            mcfd = c_int(-1)
            if rank == 0:
                rc = lib.ddpb_comm_mc_create(handle, ctypes.byref(mcfd))
                ok, why = rc == 0, (f"mc_create: {nat.error_string(rc)}" if rc else "")
            mine = [mcfd.value] if (rank == 0 and ok) else []
            got = exchange_fds(rank, world_size, mine, tok[0] + "c")
            for f in mine:
                os.close(f)
            for lst in got:
                received += lst
            if rank != 0:
                if len(got[0]) != 1:
                    ok, why = False, "rank 0 could not create the multicast object"
                else:
                    rc = lib.ddpb_comm_mc_import(handle, got[0][0])
                    ok, why = rc == 0, (f"mc_import: {nat.error_string(rc)}" if rc else "")
            if not _all_ok(ok, world_size):
                return give_up(why)
            # stage 3: every device joins the team; binding blocks until all have, so agree first
            rc = lib.ddpb_comm_mc_add_device(handle)
            ok, why = rc == 0, (f"mc_add_device: {nat.error_string(rc)}" if rc else "")
            if not _all_ok(ok, world_size):
                return give_up(why)
            rc = lib.ddpb_comm_mc_bind(handle)
            ok, why = rc == 0, (f"mc_bind: {nat.error_string(rc)}" if rc else "")
            if not _all_ok(ok, world_size):
                return give_up(why)
            nat.check(lib.ddpb_comm_set_connected(handle), "comm_set_connected")
            dist.barrier()
            self.multicast = bool(lib.ddpb_comm_has_multicast(handle))
            return handle
        finally:
            for f in received:
                try:
                    os.close(f)
                except OSError:
                    pass

    # ------------------------------------------------------------------ memory
    def tensor(self, offset_bytes: int, numel: int, dtype=torch.float32) -> torch.Tensor:
        """A torch view of ``numel`` elements of the local buffer at a 16-byte aligned offset."""
        assert offset_bytes % 16 == 0
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        assert offset_bytes + nbytes <= self.payload_bytes
        return self.buffer[offset_bytes:offset_bytes + nbytes].view(dtype)

    # ------------------------------------------------------------------ collectives
    def allreduce_sgd(self, grad_off: int, param_off: int, n_floats: int, lr_dev: torch.Tensor,
                      algo: int = ALGO_AUTO, slot: int = 0, stream=None) -> None:
        """params[slice] -= lr * mean_over_ranks(grads[slice]) in one kernel."""
        nat.check(self.lib.ddpb_allreduce_sgd(self._h, grad_off, param_off, n_floats,
                                              nat.ptr(lr_dev), algo, slot, nat.stream_ptr(stream)),
                  "allreduce_sgd")

    def reduce_allreduce_sgd(self, partials: torch.Tensor, nparts: int, n_floats: int, grad_off: int,
                             param_off: int, lr_dev: torch.Tensor, algo: int = ALGO_AUTO,
                             layout: int = LAYOUT_IDENTITY, slot: int = 0, stream=None,
                             pack: torch.Tensor | None = None, bias_parts=None) -> None:
        """One kernel: grads[slice] = sum of ``nparts`` local partial vectors (fixed order), then
        params[slice] -= lr * mean_over_ranks(grads[slice]).  ``algo``: ALGO_PUSH (default; every
        rank stores flag-carrying words straight into its peers' staging lines) or ALGO_ONESHOT
        (flag round, then pull).  ``layout`` LAYOUT_CONV3X3_WGRAD: the partials are the conv2
        weight-gradient accumulator images (``n_floats`` = CONV3X3_WGRAD_IMAGE) and the slice is
        net.2.weight followed by net.2.bias; with ``pack`` (the split bf16 tile images of
        net.2.weight, native.conv3x3_pack_split_bytes) the same kernel also rewrites those tiles from
        the updated weights; with ``bias_parts`` = (tensor, nsplit, nbx, H*W) from
        native.linear_bwd_f32 the net.2.bias gradient is their exact-fp32 sum instead of the
        accumulator images' bias column."""
        if pack is not None or bias_parts is not None:
            bp, ns, nb, hw = bias_parts if bias_parts is not None else (None, 0, 0, 0)
            nat.check(self.lib.ddpb_reduce_allreduce_sgd_pack(
                self._h, nat.ptr(partials), nparts, n_floats, grad_off, param_off, nat.ptr(lr_dev), algo,
                layout, slot, nat.ptr(pack), nat.ptr(bp), ns, nb, hw, nat.stream_ptr(stream)),
                "reduce_allreduce_sgd_pack")
            return
        nat.check(self.lib.ddpb_reduce_allreduce_sgd(self._h, nat.ptr(partials), nparts, n_floats,
                                                     grad_off, param_off, nat.ptr(lr_dev), algo, layout,
                                                     slot, nat.stream_ptr(stream)),
                  "reduce_allreduce_sgd")

    def allreduce(self, in_off: int, out_off: int, nbytes: int, dtype=torch.float32,
                  algo: int = ALGO_AUTO, slot: int = 0, stream=None) -> None:
        nat.check(self.lib.ddpb_allreduce(self._h, in_off, out_off, nbytes, DTYPE_CODES[dtype], algo,
                                          slot, nat.stream_ptr(stream)), "allreduce")

    def broadcast(self, off: int, nbytes: int, root: int = 0, slot: int = 0, stream=None) -> None:
        nat.check(self.lib.ddpb_broadcast(self._h, off, nbytes, root, slot, nat.stream_ptr(stream)),
                  "broadcast")

    def status(self) -> int:
        s = c_uint(0)
        nat.check(self.lib.ddpb_comm_status(self._h, ctypes.byref(s)), "comm_status")
        return s.value

    def close(self) -> None:
        if self._h is not None:
            self.buffer = None
            self.lib.ddpb_comm_destroy(self._h)
            self._h = None
