"""Symmetric-memory communicator (Python side).

Owns this rank's peer-mapped buffer (native: csrc/comm.cu) and exposes slices of it as torch
tensors, plus the in-kernel-flag collectives that replace what the reference gets from
ProcessGroupNCCL + the c10d Reducer (reference utils.py:7-11, train_ddp.py:34,199;
SURVEY.md §2.5 X13, X16, §5.9).  torch.distributed is used once, as plumbing, to exchange the
64-byte CUDA IPC handles.

All ranks must issue the same sequence of collective calls per flag slot.
"""
from __future__ import annotations

import ctypes
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
    ("ddpb_allreduce_sgd", I, [P, SZ, SZ, LL, P, I, I, P]),
    ("ddpb_reduce_allreduce_sgd", I, [P, P, I, LL, SZ, SZ, P, I, I, P]),
    ("ddpb_allreduce", I, [P, SZ, SZ, LL, I, I, I, P]),
    ("ddpb_broadcast", I, [P, SZ, LL, I, I, P]),
):
    nat.declare(_name, _res, _args)

ALGO_AUTO, ALGO_ONESHOT, ALGO_TWOSHOT, ALGO_PUSH = 0, 1, 2, 3
DTYPE_CODES = {torch.float32: 0, torch.bfloat16: 1}


class _RawCudaBuffer:
    """Adapter so torch can wrap a raw device pointer without copying."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2, "strides": None,
        }


class SymmetricComm:
    def __init__(self, rank: int, world_size: int, payload_bytes: int, timeout_ms: int | None = None):
        lib = nat.lib()
        self.lib = lib
        self.rank, self.world_size = rank, world_size
        handle = c_void_p()
        nat.check(lib.ddpb_comm_create(rank, world_size, payload_bytes, ctypes.byref(handle)),
                  "comm_create")
        self._h = handle
        if timeout_ms is not None:
            lib.ddpb_comm_set_timeout_ms(self._h, int(timeout_ms))
        if world_size > 1:
            hb = lib.ddpb_comm_handle_bytes()
            mine = ctypes.create_string_buffer(hb)
            nat.check(lib.ddpb_comm_get_handle(self._h, mine), "comm_get_handle")
            gathered = [None] * world_size
            dist.all_gather_object(gathered, bytes(mine.raw))
            blob = b"".join(gathered)
            nat.check(lib.ddpb_comm_connect(self._h, blob), "comm_connect (CUDA IPC open)")
            dist.barrier()
        self.payload_bytes = int(lib.ddpb_comm_payload_bytes(self._h))
        self._base_ptr = int(lib.ddpb_comm_local_ptr(self._h))
        self._keep = _RawCudaBuffer(self._base_ptr, self.payload_bytes)
        self.buffer = torch.as_tensor(self._keep, device=torch.device("cuda", torch.cuda.current_device()))

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
                             slot: int = 0, stream=None) -> None:
        """One kernel: grads[slice] = sum of ``nparts`` local partial vectors (fixed order), then
        params[slice] -= lr * mean_over_ranks(grads[slice]).  ``algo``: ALGO_PUSH (default; every
        rank stores flag-carrying words straight into its peers' staging lines) or ALGO_ONESHOT
        (flag round, then pull)."""
        nat.check(self.lib.ddpb_reduce_allreduce_sgd(self._h, nat.ptr(partials), nparts, n_floats,
                                                     grad_off, param_off, nat.ptr(lr_dev), algo, slot,
                                                     nat.stream_ptr(stream)), "reduce_allreduce_sgd")

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
