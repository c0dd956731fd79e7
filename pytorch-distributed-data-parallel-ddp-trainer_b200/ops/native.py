"""ctypes binding of the native sm_100a library (``lib/libddp_b200.so``; C ABI in
``include/ddp_b200.h``; sources in ``csrc/``).

The reference has no native layer — every op below replaces a library call it delegates to
PyTorch (cuDNN / cuBLAS / ATen / c10d).  There is deliberately no Python
fallback here: on a GPU box a missing or stale library is an error, not a silent downgrade.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_longlong, c_size_t, c_void_p, POINTER

import torch

_PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(_PKG_DIR, "lib", "libddp_b200.so")

_lib = None

P = c_void_p
I = c_int
LL = c_longlong

# name -> (restype, argtypes)
_SIGNATURES = {
    "ddpb_error_string": (c_char_p, [I]),
    "ddpb_num_sms": (I, []),
    "ddpb_plane_rows": (LL, [I, I, I]),
    "ddpb_guard_rows": (I, []),
    "ddpb_tile_rows": (I, []),
    "ddpb_gather_u8": (I, [P, P, P, P, P, I, I, P]),
    "ddpb_conv1_fwd": (I, [P, P, P, P, I, I, I, I, I, I, P]),
    "ddpb_conv1_wgrad_scratch_bytes": (c_size_t, [I, I, I, I, I]),
    "ddpb_conv1_wgrad": (I, [P, P, P, P, I, I, I, I, I, I, P]),
    "ddpb_conv1_fwd_tc": (I, [P, P, P, P, I, I, I, I, I, I, P]),
    "ddpb_conv1_wgrad_tc_scratch_bytes": (c_size_t, [I, I]),
    "ddpb_conv1_wgrad_tc": (I, [P, P, P, P, I, I, I, I, I, I, P]),
    "ddpb_conv1_wgrad_tc_partials": (I, [P, P, P, POINTER(c_int), I, I, I, I, I, I, P]),
    "ddpb_conv3x3_fwd": (I, [P, P, P, P, I, I, I, I, I, I, P]),
    "ddpb_conv3x3_dgrad": (I, [P, P, P, P, I, I, I, I, I, I, P]),
    "ddpb_conv3x3_wgrad_scratch_bytes": (c_size_t, [I, I]),
    "ddpb_conv3x3_wgrad": (I, [P, P, P, P, P, I, I, I, I, I, I, P]),
    "ddpb_conv3x3_wgrad_partials": (I, [P, P, P, POINTER(c_int), I, I, I, I, I, I, P]),
    "ddpb_linear_fwd_scratch_bytes": (c_size_t, [I, I, I, I, I]),
    "ddpb_linear_fwd": (I, [P, P, P, POINTER(c_int), I, I, I, I, I, I, P]),
    "ddpb_softmax_xent": (I, [P, I, P, P, P, P, P, P, I, I, P]),
    "ddpb_softmax_xent_packed": (I, [P, I, P, P, P, P, P, P, I, I, P, P]),
    "ddpb_linear_pack_bytes": (c_size_t, [I, I, I, I]),
    "ddpb_linear_pack": (I, [P, P, I, I, I, I, P]),
    "ddpb_linear_fwd_tc_scratch_bytes": (c_size_t, [I, I, I, I, I]),
    "ddpb_linear_fwd_tc": (I, [P, P, P, POINTER(c_int), I, I, I, I, I, I, P]),
    "ddpb_linear_bwd_tc": (I, [P, P, P, P, P, I, I, I, I, I, I, P]),
    "ddpb_set_pdl": (None, [I]),
    "ddpb_get_pdl": (I, []),
    "ddpb_linear_bwd_scratch_bytes": (c_size_t, [I, I, I, I, I]),
    "ddpb_linear_bwd": (I, [P, P, P, P, P, P, I, I, I, I, I, I, P]),
    "ddpb_sgd": (I, [P, P, P, LL, P]),
    # split planes (x = hi + lo bf16, "bf16x3" GEMMs) + fp32 linear layer: the reference-precision mode
    "ddpb_conv1_fwd_tc_split": (I, [P, P, P, P, I, I, I, I, I, I, P]),
    "ddpb_conv1_wgrad_tc_split": (I, [P, P, P, P, I, I, I, I, I, I, P]),
    "ddpb_conv1_wgrad_tc_split_partials": (I, [P, P, P, POINTER(c_int), I, I, I, I, I, I, P]),
    "ddpb_conv3x3_pack_split_bytes": (c_size_t, [I, I]),
    "ddpb_conv3x3_pack_split": (I, [P, P, I, I, P]),
    "ddpb_conv3x3_fwd_split": (I, [P, P, P, P, I, I, I, I, I, I, I, P]),
    "ddpb_conv3x3_dgrad_split": (I, [P, P, P, P, I, I, I, I, I, I, I, P]),
    "ddpb_conv3x3_wgrad_split": (I, [P, P, P, P, P, P, I, I, I, I, I, I, I, I, P]),
    "ddpb_conv3x3_wgrad_split_partials": (I, [P, P, P, POINTER(c_int), I, I, I, I, I, I, P]),
    "ddpb_linear_fwd_f32_scratch_bytes": (c_size_t, [I, I, I, I, I]),
    "ddpb_linear_fwd_f32": (I, [P, P, P, POINTER(c_int), I, I, I, I, I, I, P]),
    "ddpb_linear_fwd_f32_xent_max_batch": (I, []),
    "ddpb_linear_fwd_f32_xent": (I, [P, P, P, P, P, P, P, P, P, I, I, I, I, I, I, P]),
    "ddpb_linear_bwd_f32_bias_part_bytes": (c_size_t, [I, I, I, I]),
    "ddpb_linear_bwd_f32": (I, [P, P, P, P, P, P, I, I, I, I, I, I, I, P, POINTER(c_int), POINTER(c_int), P]),
    "ddpb_linear_bwd_f32_mode": (I, [P, P, P, P, P, P, I, I, I, I, I, I, I, P, POINTER(c_int), POINTER(c_int), I, P]),
    "ddpb_conv1_wgrad_f32_scratch_bytes": (c_size_t, [I, I]),
    "ddpb_conv1_wgrad_f32": (I, [P, P, P, P, I, I, I, I, I, I, P]),
    "ddpb_conv1_wgrad_f32_partials": (I, [P, P, P, POINTER(c_int), I, I, I, I, I, I, P]),
    "ddpb_zero_tail_rows": (I, [P, I, I, I, I, I, P]),
    "ddpb_reduce_partials": (I, [P, I, LL, P, P]),
    # fp32 planes, tf32 tensor-core operands: forward / data-gradient building blocks
    "ddpb_conv1_fwd_tc_f32": (I, [P, P, P, P, I, I, I, I, I, I, P]),
    "ddpb_conv3x3_pack_tf32_bytes": (c_size_t, [I, I]),
    "ddpb_conv3x3_pack_tf32": (I, [P, P, I, I, P]),
    "ddpb_conv3x3_fwd_tf32": (I, [P, P, P, P, I, I, I, I, I, I, I, P]),
    "ddpb_conv3x3_dgrad_tf32": (I, [P, P, P, P, I, I, I, I, I, I, I, P]),
    # fully connected layers of any shape (csrc/gemm_tc.cu) + fused ReLU/max-pool
    "ddpb_gemm_pad": (I, []),
    "ddpb_pack_rows_split_bytes": (c_size_t, [I, I]),
    "ddpb_pack_rows_split": (I, [P, P, LL, I, I, P, P]),
    "ddpb_gemm_split": (I, [P, I, I, I, P, I, I, I, I, I, I, P, LL, I, I, LL, P]),
    "ddpb_gemm_split_suggest_splits": (I, [I, I, I]),
    "ddpb_splitk_reduce": (I, [P, I, LL, I, I, LL, P, I, P, LL, P]),
    "ddpb_colsum": (I, [P, P, I, I, LL, P, I, P]),
    "ddpb_relu_maxpool2_fwd": (I, [P, P, P, P, I, I, I, I, P]),
    "ddpb_relu_maxpool2_bwd": (I, [P, P, P, I, I, I, I, P]),
    # channels-last kernels of the widened CNN (csrc/wide_ops.cu)
    "ddpb_conv3x3_small_cin_fwd": (I, [P, P, P, P, I, I, I, I, I, P]),
    "ddpb_conv3x3_small_cin_wgrad_parts": (I, [I, I]),
    "ddpb_conv3x3_small_cin_wgrad": (I, [P, P, P, P, P, I, I, I, I, I, P]),
    "ddpb_relu_maxpool2_nhwc_fwd": (I, [P, P, P, P, I, I, I, I, P]),
    "ddpb_relu_maxpool2_nhwc_bwd": (I, [P, P, P, I, I, I, I, P]),
    "ddpb_small_linear_max_out": (I, []),
    "ddpb_small_linear_fwd": (I, [P, P, P, P, I, I, I, P]),
    "ddpb_small_linear_bwd_scratch_floats": (c_size_t, [I, I, I]),
    "ddpb_small_linear_bwd": (I, [P, P, P, P, P, P, P, I, I, I, P]),
}

#: execution precisions of the step (models/simple_cnn.py): "bf16x3" = the reference's numerics
#: (fp32-equivalent storage, >= TF32 operand precision for the convolutions, fp32 linear layer),
#: "bf16" = fast mode
PRECISIONS = ("bf16x3", "bf16")
#: plane storage formats (csrc/layout.cuh)
PLANE_FORMATS = ("bf16", "split", "f32")


class NativeError(RuntimeError):
    pass


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib() -> ctypes.CDLL:
    """Load (once) and return the native library; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NativeError(
                f"{LIB_PATH} is missing: run `python build.py` (or __graft_entry__.build()) first")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name, None)
            if fn is None:
                continue  # optional symbols are declared where they are used
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def declare(name: str, restype, argtypes) -> None:
    """Register the signature of an entry point defined by another module of this package."""
    _SIGNATURES[name] = (restype, argtypes)
    if _lib is not None:
        fn = getattr(_lib, name)
        fn.restype = restype
        fn.argtypes = argtypes


def check(code: int, what: str = "") -> None:
    if code != 0:
        msg = lib().ddpb_error_string(int(code))
        raise NativeError(f"{what or 'native call'} failed: [{code}] {msg.decode() if msg else '?'}")


def error_string(code: int) -> str:
    msg = lib().ddpb_error_string(int(code))
    return f"[{code}] {msg.decode() if msg else '?'}"


def ptr(t) -> int | None:
    """Device/host pointer of a tensor (None passes NULL)."""
    if t is None:
        return None
    return t.data_ptr()


def stream_ptr(stream=None) -> int:
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


# ----------------------------------------------------------------------------- geometry
def plane_rows(batch_cap: int, H: int, W: int) -> int:
    return int(lib().ddpb_plane_rows(batch_cap, H, W))


def guard_rows() -> int:
    return int(lib().ddpb_guard_rows())


def alloc_planes(channels: int, batch_cap: int, H: int, W: int, device="cuda",
                 fmt: str = "bf16") -> torch.Tensor:
    """Zero-filled activation buffer (16-byte rows in every format, csrc/layout.cuh):
    "bf16"  [C/8, plane_rows, 8] bf16;
    "split" [2*C/8, plane_rows, 8] bf16 — C/8 hi planes then C/8 lo planes, value = hi + lo;
    "f32"   [C/4, plane_rows, 4] fp32."""
    assert fmt in PLANE_FORMATS
    rows = plane_rows(batch_cap, H, W)
    if fmt == "f32":
        assert channels % 4 == 0
        return torch.zeros(channels // 4, rows, 4, dtype=torch.float32, device=device)
    assert channels % 8 == 0
    sets = 2 if fmt == "split" else 1
    return torch.zeros(sets * channels // 8, rows, 8, dtype=torch.bfloat16, device=device)


def scratch(nbytes: int, device="cuda") -> torch.Tensor:
    return torch.empty(max(int(nbytes), 16) // 4 + 4, dtype=torch.float32, device=device)


# ----------------------------------------------------------------------------- op wrappers
def gather_u8(images, labels, idx, x, y, stream=None):
    batch = idx.numel()
    pix = images[0].numel()
    check(lib().ddpb_gather_u8(ptr(images), ptr(labels), ptr(idx), ptr(x), ptr(y), batch, pix,
                               stream_ptr(stream)), "gather_u8")


def conv1_fwd(x, w, b, out, batch, batch_cap, H, W, stream=None):
    check(lib().ddpb_conv1_fwd(ptr(x), ptr(w), ptr(b), ptr(out), w.shape[1], w.shape[0], batch,
                               batch_cap, H, W, stream_ptr(stream)), "conv1_fwd")


def conv1_wgrad_scratch_bytes(cin, cout, batch_cap, H, W) -> int:
    return int(lib().ddpb_conv1_wgrad_scratch_bytes(cin, cout, batch_cap, H, W))


def conv1_wgrad(dact, x, dw_db, scr, cin, cout, batch, batch_cap, H, W, stream=None):
    check(lib().ddpb_conv1_wgrad(ptr(dact), ptr(x), ptr(dw_db), ptr(scr), cin, cout, batch,
                                 batch_cap, H, W, stream_ptr(stream)), "conv1_wgrad")


def conv1_fwd_tc(x, w, b, out, batch, batch_cap, H, W, stream=None):
    check(lib().ddpb_conv1_fwd_tc(ptr(x), ptr(w), ptr(b), ptr(out), w.shape[1], w.shape[0], batch,
                                  batch_cap, H, W, stream_ptr(stream)), "conv1_fwd_tc")


def conv1_wgrad_tc_scratch_bytes(cin, cout) -> int:
    return int(lib().ddpb_conv1_wgrad_tc_scratch_bytes(cin, cout))


def conv1_wgrad_tc(dact, x, dw_db, scr, cin, cout, batch, batch_cap, H, W, stream=None):
    check(lib().ddpb_conv1_wgrad_tc(ptr(dact), ptr(x), ptr(dw_db), ptr(scr), cin, cout, batch,
                                    batch_cap, H, W, stream_ptr(stream)), "conv1_wgrad_tc")


def conv1_wgrad_tc_partials(dact, x, scr, cin, cout, batch, batch_cap, H, W, stream=None) -> int:
    """Per-CTA partial gradients only (scr[nparts][cout*cin*9 + cout]); returns nparts."""
    n = c_int(0)
    check(lib().ddpb_conv1_wgrad_tc_partials(ptr(dact), ptr(x), ptr(scr), ctypes.byref(n), cin, cout,
                                             batch, batch_cap, H, W, stream_ptr(stream)),
          "conv1_wgrad_tc_partials")
    return n.value


def conv3x3_fwd(inp, w, bias, out, batch, batch_cap, H, W, stream=None):
    check(lib().ddpb_conv3x3_fwd(ptr(inp), ptr(w), ptr(bias), ptr(out), w.shape[1], w.shape[0],
                                 batch, batch_cap, H, W, stream_ptr(stream)), "conv3x3_fwd")


def conv3x3_dgrad(dout, w, mask, din, batch, batch_cap, H, W, stream=None):
    check(lib().ddpb_conv3x3_dgrad(ptr(dout), ptr(w), ptr(mask), ptr(din), w.shape[1], w.shape[0],
                                   batch, batch_cap, H, W, stream_ptr(stream)), "conv3x3_dgrad")


def conv3x3_wgrad_scratch_bytes(cin, cout) -> int:
    return int(lib().ddpb_conv3x3_wgrad_scratch_bytes(cin, cout))


def conv3x3_wgrad(dout, inp, dw, db, scr, cin, cout, batch, batch_cap, H, W, stream=None):
    check(lib().ddpb_conv3x3_wgrad(ptr(dout), ptr(inp), ptr(dw), ptr(db), ptr(scr), cin, cout,
                                   batch, batch_cap, H, W, stream_ptr(stream)), "conv3x3_wgrad")


def conv3x3_wgrad_partials(dout, inp, scr, cin, cout, batch, batch_cap, H, W, stream=None) -> int:
    """Per-CTA accumulator images only (scr[nparts][cout*(9*cin+8)]); returns nparts.  Consumed by
    SymmetricComm.reduce_allreduce_sgd(..., layout=LAYOUT_CONV3X3_WGRAD)."""
    n = c_int(0)
    check(lib().ddpb_conv3x3_wgrad_partials(ptr(dout), ptr(inp), ptr(scr), ctypes.byref(n), cin, cout,
                                            batch, batch_cap, H, W, stream_ptr(stream)),
          "conv3x3_wgrad_partials")
    return n.value


def linear_fwd_scratch_bytes(nout, C, batch_cap, H, W) -> int:
    return int(lib().ddpb_linear_fwd_scratch_bytes(nout, C, batch_cap, H, W))


def linear_fwd(act, Wfl, partial, nout, C, batch, batch_cap, H, W, stream=None) -> int:
    """Returns S, the number of partial slices written."""
    s = c_int(0)
    check(lib().ddpb_linear_fwd(ptr(act), ptr(Wfl), ptr(partial), ctypes.byref(s), nout, C, batch,
                                batch_cap, H, W, stream_ptr(stream)), "linear_fwd")
    return s.value


def linear_pack_bytes(nout, C, H, W) -> int:
    return int(lib().ddpb_linear_pack_bytes(nout, C, H, W))


def linear_pack(Wfl, packed, nout, C, H, W, stream=None):
    """fp32 linear weights -> bf16 UMMA tiles (hi+lo forward tile, backward tile)."""
    check(lib().ddpb_linear_pack(ptr(Wfl), ptr(packed), nout, C, H, W, stream_ptr(stream)), "linear_pack")


def linear_fwd_tc_scratch_bytes(nout, C, batch_cap, H, W) -> int:
    return int(lib().ddpb_linear_fwd_tc_scratch_bytes(nout, C, batch_cap, H, W))


def linear_fwd_tc(act, packed, partial, nout, C, batch, batch_cap, H, W, stream=None) -> int:
    """Tensor-core linear forward; returns S, the number of partial slices written."""
    s = c_int(0)
    check(lib().ddpb_linear_fwd_tc(ptr(act), ptr(packed), ptr(partial), ctypes.byref(s), nout, C, batch,
                                   batch_cap, H, W, stream_ptr(stream)), "linear_fwd_tc")
    return s.value


def linear_bwd_tc(act, dl_packed, packed, dact, dW, nout, C, batch, batch_cap, H, W, stream=None):
    """Tensor-core linear backward: dact = relu'(act) * dlogits.W, dW = dlogits^T.act (no scratch).
    ``dl_packed`` is the tile buffer written by softmax_xent(dl_packed=...)."""
    check(lib().ddpb_linear_bwd_tc(ptr(act), ptr(dl_packed), ptr(packed), ptr(dact), ptr(dW), nout, C,
                                   batch, batch_cap, H, W, stream_ptr(stream)), "linear_bwd_tc")


def dl_packed_bytes(batch_cap: int) -> int:
    """Size of the bf16 hi/lo dlogits tile buffer softmax_xent(dl_packed=...) writes."""
    return (batch_cap + 127) // 128 * 128 * 64


def softmax_xent(partial, S, bias, labels, logits, dlogits, dbias, loss, nout, batch, stream=None,
                 dl_packed=None):
    if dl_packed is None:
        check(lib().ddpb_softmax_xent(ptr(partial), S, ptr(bias), ptr(labels), ptr(logits),
                                      ptr(dlogits), ptr(dbias), ptr(loss), nout, batch,
                                      stream_ptr(stream)), "softmax_xent")
    else:
        check(lib().ddpb_softmax_xent_packed(ptr(partial), S, ptr(bias), ptr(labels), ptr(logits),
                                             ptr(dlogits), ptr(dbias), ptr(loss), nout, batch,
                                             ptr(dl_packed), stream_ptr(stream)), "softmax_xent_packed")


def linear_bwd_scratch_bytes(nout, C, batch_cap, H, W) -> int:
    return int(lib().ddpb_linear_bwd_scratch_bytes(nout, C, batch_cap, H, W))


def linear_bwd(act, dlogits, Wfl, dact, dW, scr, nout, C, batch, batch_cap, H, W, stream=None):
    check(lib().ddpb_linear_bwd(ptr(act), ptr(dlogits), ptr(Wfl), ptr(dact), ptr(dW), ptr(scr),
                                nout, C, batch, batch_cap, H, W, stream_ptr(stream)), "linear_bwd")


def sgd(p, grad, lr_dev, stream=None):
    check(lib().ddpb_sgd(ptr(p), ptr(grad), ptr(lr_dev), p.numel(), stream_ptr(stream)), "sgd")


# ----------------------------------------------------------------------------- split planes / fp32 linear
def conv1_fwd_tc_split(x, w, b, out, batch, batch_cap, H, W, stream=None):
    """conv1 + bias + ReLU into split planes (hi + lo bf16)."""
    check(lib().ddpb_conv1_fwd_tc_split(ptr(x), ptr(w), ptr(b), ptr(out), w.shape[1], w.shape[0], batch,
                                        batch_cap, H, W, stream_ptr(stream)), "conv1_fwd_tc_split")


def conv1_wgrad_tc_split(dact, x, dw_db, scr, cin, cout, batch, batch_cap, H, W, stream=None):
    check(lib().ddpb_conv1_wgrad_tc_split(ptr(dact), ptr(x), ptr(dw_db), ptr(scr), cin, cout, batch,
                                          batch_cap, H, W, stream_ptr(stream)), "conv1_wgrad_tc_split")


def conv1_wgrad_tc_split_partials(dact, x, scr, cin, cout, batch, batch_cap, H, W, stream=None) -> int:
    n = c_int(0)
    check(lib().ddpb_conv1_wgrad_tc_split_partials(ptr(dact), ptr(x), ptr(scr), ctypes.byref(n), cin, cout,
                                                   batch, batch_cap, H, W, stream_ptr(stream)),
          "conv1_wgrad_tc_split_partials")
    return n.value


def conv3x3_pack_split_bytes(cin, cout) -> int:
    return int(lib().ddpb_conv3x3_pack_split_bytes(cin, cout))


def conv3x3_pack_split(w, packed, stream=None):
    """fp32 conv weight [cout,cin,3,3] -> split bf16 tile images [fwd hi|fwd lo|dgrad hi|dgrad lo]."""
    check(lib().ddpb_conv3x3_pack_split(ptr(w), ptr(packed), w.shape[1], w.shape[0], stream_ptr(stream)),
          "conv3x3_pack_split")


def conv3x3_fwd_split(inp, packed, bias, out, cin, cout, batch, batch_cap, H, W, stream=None,
                      out_f32: bool = True):
    """Split-plane input; ``out`` = fp32 planes (default: the consumer is the fp32 linear layer) or
    split planes."""
    check(lib().ddpb_conv3x3_fwd_split(ptr(inp), ptr(packed), ptr(bias), ptr(out), cin, cout, batch,
                                       batch_cap, H, W, int(out_f32), stream_ptr(stream)),
          "conv3x3_fwd_split")


def conv3x3_dgrad_split(dout, packed, mask, din, cin, cout, batch, batch_cap, H, W, stream=None,
                        out_f32: bool = False):
    """``din`` = split planes, or unrounded fp32 planes with ``out_f32`` (for the exact-fp32 conv1
    weight gradient)."""
    check(lib().ddpb_conv3x3_dgrad_split(ptr(dout), ptr(packed), ptr(mask), ptr(din), cin, cout, batch,
                                         batch_cap, H, W, int(out_f32), stream_ptr(stream)),
          "conv3x3_dgrad_split")


def conv3x3_wgrad_split(dout, inp, dw, db, scr, cin, cout, batch, batch_cap, H, W, stream=None,
                        bias_parts=None):
    """``bias_parts`` = (tensor, nsplit, nbx) from linear_bwd_f32: db becomes their exact-fp32 sum."""
    bp, ns, nb = bias_parts if bias_parts is not None else (None, 0, 0)
    check(lib().ddpb_conv3x3_wgrad_split(ptr(dout), ptr(inp), ptr(dw), ptr(db), ptr(scr), ptr(bp), ns, nb,
                                         cin, cout, batch, batch_cap, H, W, stream_ptr(stream)),
          "conv3x3_wgrad_split")


def conv3x3_wgrad_split_partials(dout, inp, scr, cin, cout, batch, batch_cap, H, W, stream=None) -> int:
    n = c_int(0)
    check(lib().ddpb_conv3x3_wgrad_split_partials(ptr(dout), ptr(inp), ptr(scr), ctypes.byref(n), cin,
                                                  cout, batch, batch_cap, H, W, stream_ptr(stream)),
          "conv3x3_wgrad_split_partials")
    return n.value


def linear_fwd_f32_scratch_bytes(nout, C, batch_cap, H, W) -> int:
    return int(lib().ddpb_linear_fwd_f32_scratch_bytes(nout, C, batch_cap, H, W))


def linear_fwd_f32(act, Wfl, partial, nout, C, batch, batch_cap, H, W, stream=None) -> int:
    """fp32 linear forward on fp32 planes; returns S, the number of partial slices written."""
    s = c_int(0)
    check(lib().ddpb_linear_fwd_f32(ptr(act), ptr(Wfl), ptr(partial), ctypes.byref(s), nout, C, batch,
                                    batch_cap, H, W, stream_ptr(stream)), "linear_fwd_f32")
    return s.value


def linear_fwd_f32_xent_max_batch() -> int:
    return int(lib().ddpb_linear_fwd_f32_xent_max_batch())


def linear_fwd_f32_xent(act, Wfl, partial, bias, labels, logits, dlogits, dbias, loss, nout, C, batch,
                        batch_cap, H, W, stream=None):
    """fp32 linear forward with the softmax cross-entropy loss, dlogits and dbias computed in the
    tail of the same kernel (small batches: one launch less on the step's critical path)."""
    check(lib().ddpb_linear_fwd_f32_xent(ptr(act), ptr(Wfl), ptr(partial), ptr(bias), ptr(labels),
                                         ptr(logits), ptr(dlogits), ptr(dbias), ptr(loss), nout, C, batch,
                                         batch_cap, H, W, stream_ptr(stream)), "linear_fwd_f32_xent")


def linear_bwd_f32_bias_part_bytes(C, batch_cap, H, W) -> int:
    return int(lib().ddpb_linear_bwd_f32_bias_part_bytes(C, batch_cap, H, W))


def linear_bwd_f32(act, dlogits, Wfl, dact, dW, scr, nout, C, batch, batch_cap, H, W, stream=None,
                   split_out: bool = True, bias_part=None, mode: int = 0):
    """fp32 linear backward on fp32 activation planes; ``dact`` = split planes (default) or fp32
    planes rounded to tf32.  With ``bias_part`` (linear_bwd_f32_bias_part_bytes) the kernel also leaves
    the exact-fp32 per-channel sums of dact there; returns the extents (nsplit, nbx) of this launch.
    ``mode`` 1 = the dX half only (dact + bias parts), 2 = the dW half only, 0 = both."""
    ns, nb = c_int(0), c_int(0)
    check(lib().ddpb_linear_bwd_f32_mode(ptr(act), ptr(dlogits), ptr(Wfl), ptr(dact), ptr(dW), ptr(scr),
                                         nout, C, batch, batch_cap, H, W, int(split_out), ptr(bias_part),
                                         ctypes.byref(ns), ctypes.byref(nb), int(mode), stream_ptr(stream)),
          "linear_bwd_f32")
    return ns.value, nb.value


def reduce_partials(part, nparts: int, n: int, out, stream=None):
    """out[i] = sum_p part[p*n + i], fixed order."""
    check(lib().ddpb_reduce_partials(ptr(part), nparts, n, ptr(out), stream_ptr(stream)), "reduce_partials")


def conv1_wgrad_f32_scratch_bytes(cin, cout) -> int:
    return int(lib().ddpb_conv1_wgrad_f32_scratch_bytes(cin, cout))


def conv1_wgrad_f32(dact, x, dw_db, scr, cin, cout, batch, batch_cap, H, W, stream=None):
    """Exact-fp32 conv1 weight + bias gradient from fp32 planes (CUDA cores)."""
    check(lib().ddpb_conv1_wgrad_f32(ptr(dact), ptr(x), ptr(dw_db), ptr(scr), cin, cout, batch,
                                     batch_cap, H, W, stream_ptr(stream)), "conv1_wgrad_f32")


def conv1_wgrad_f32_partials(dact, x, scr, cin, cout, batch, batch_cap, H, W, stream=None) -> int:
    n = c_int(0)
    check(lib().ddpb_conv1_wgrad_f32_partials(ptr(dact), ptr(x), ptr(scr), ctypes.byref(n), cin, cout,
                                              batch, batch_cap, H, W, stream_ptr(stream)),
          "conv1_wgrad_f32_partials")
    return n.value


# ----------------------------------------------------------------------------- fp32 planes / tf32 blocks
def conv1_fwd_tc_f32(x, w, b, out, batch, batch_cap, H, W, stream=None):
    """conv1 + bias + ReLU into fp32 planes (values rounded to tf32)."""
    check(lib().ddpb_conv1_fwd_tc_f32(ptr(x), ptr(w), ptr(b), ptr(out), w.shape[1], w.shape[0], batch,
                                      batch_cap, H, W, stream_ptr(stream)), "conv1_fwd_tc_f32")


def conv3x3_pack_tf32_bytes(cin, cout) -> int:
    return int(lib().ddpb_conv3x3_pack_tf32_bytes(cin, cout))


def conv3x3_pack_tf32(w, packed, stream=None):
    """fp32 conv weight [cout,cin,3,3] -> tf32 tile images (forward + data-gradient) in ``packed``."""
    check(lib().ddpb_conv3x3_pack_tf32(ptr(w), ptr(packed), w.shape[1], w.shape[0], stream_ptr(stream)),
          "conv3x3_pack_tf32")


def conv3x3_fwd_tf32(inp, packed, bias, out, cin, cout, batch, batch_cap, H, W, stream=None,
                     round_out: bool = False):
    check(lib().ddpb_conv3x3_fwd_tf32(ptr(inp), ptr(packed), ptr(bias), ptr(out), cin, cout, batch,
                                      batch_cap, H, W, int(round_out), stream_ptr(stream)),
          "conv3x3_fwd_tf32")


def conv3x3_dgrad_tf32(dout, packed, mask, din, cin, cout, batch, batch_cap, H, W, stream=None,
                       round_out: bool = True):
    check(lib().ddpb_conv3x3_dgrad_tf32(ptr(dout), ptr(packed), ptr(mask), ptr(din), cin, cout, batch,
                                        batch_cap, H, W, int(round_out), stream_ptr(stream)),
          "conv3x3_dgrad_tf32")


# ----------------------------------------------------------------------------- general linear layers
class PackedMatrix:
    """A fp32 matrix [rows, k] packed as split row planes (csrc/gemm_tc.cu): the operand format of
    ``gemm_split``.  ``rows`` may change from call to call up to the capacity it was created with."""

    def __init__(self, rows_cap: int, k: int, device="cuda"):
        self.rows_cap, self.k, self.rows = int(rows_cap), int(k), 0
        nbytes = int(lib().ddpb_pack_rows_split_bytes(self.rows_cap, self.k))
        self.buf = torch.zeros(nbytes // 2, dtype=torch.bfloat16, device=device)

    def pack(self, src: torch.Tensor, mask: torch.Tensor | None = None, stream=None) -> "PackedMatrix":
        assert src.dim() == 2 and src.dtype == torch.float32 and src.stride(1) == 1, "row-major fp32 matrix"
        rows, k = src.shape
        assert k == self.k and rows <= self.rows_cap, (src.shape, self.rows_cap, self.k)
        if mask is not None:
            assert mask.shape == src.shape and mask.dtype == torch.float32 and mask.stride() == src.stride()
        check(lib().ddpb_pack_rows_split(ptr(src), ptr(mask), src.stride(0), rows, k, ptr(self.buf),
                                         stream_ptr(stream)), "pack_rows_split")
        self.rows = rows
        return self


def gemm_split(a: PackedMatrix, a_mn: bool, b: PackedMatrix, b_mn: bool, M: int, N: int, K: int,
               out: torch.Tensor, ld: int, transposed: bool, k_splits: int = 1, split_stride: int = 0,
               stream=None):
    """out[m][n] (out[n][m] when ``transposed``) = sum_k A[m][k] * B[n][k] on the tensor cores (bf16x3)."""
    check(lib().ddpb_gemm_split(ptr(a.buf), a.rows_cap, a.k, int(a_mn), ptr(b.buf), b.rows_cap, b.k, int(b_mn),
                                M, N, K, ptr(out), ld, int(transposed), k_splits, split_stride,
                                stream_ptr(stream)), "gemm_split")


def gemm_suggest_splits(M: int, N: int, K: int) -> int:
    return int(lib().ddpb_gemm_split_suggest_splits(M, N, K))


def splitk_reduce(partial, nsplit, split_stride, rows, cols, ld, bias, relu, out, ld_out=None, stream=None):
    check(lib().ddpb_splitk_reduce(ptr(partial), nsplit, split_stride, rows, cols, ld, ptr(bias), int(relu),
                                   ptr(out), ld if ld_out is None else ld_out, stream_ptr(stream)),
          "splitk_reduce")


def colsum(src, mask, out, stream=None):
    """out[c] = sum_r src[r][c] where mask[r][c] > 0 (mask optional); tall matrices in two fixed-order passes."""
    rows, cols = src.shape
    splits = 1 if rows <= 1024 else min(512, (rows + 255) // 256)
    if splits == 1:
        check(lib().ddpb_colsum(ptr(src), ptr(mask), rows, cols, src.stride(0), ptr(out), 1, stream_ptr(stream)),
              "colsum")
        return
    rps = (rows + splits - 1) // splits
    splits = (rows + rps - 1) // rps
    part = torch.empty(splits, cols, dtype=torch.float32, device=src.device)
    check(lib().ddpb_colsum(ptr(src), ptr(mask), rows, cols, src.stride(0), ptr(part), splits, stream_ptr(stream)),
          "colsum")
    check(lib().ddpb_colsum(ptr(part), None, splits, cols, cols, ptr(out), 1, stream_ptr(stream)), "colsum")


def relu_maxpool2_fwd(x, bias, out, code, stream=None):
    b, c, h, w = x.shape
    check(lib().ddpb_relu_maxpool2_fwd(ptr(x), ptr(bias), ptr(out), ptr(code), b, c, h, w, stream_ptr(stream)),
          "relu_maxpool2_fwd")


def relu_maxpool2_bwd(dout, code, din, stream=None):
    b, c, h, w = din.shape
    check(lib().ddpb_relu_maxpool2_bwd(ptr(dout), ptr(code), ptr(din), b, c, h, w, stream_ptr(stream)),
          "relu_maxpool2_bwd")


# ----------------------------------------------------------------------------- channels-last kernels (widened CNN)
def conv3x3_small_cin_fwd(x, w, bias, out_nhwc, stream=None):
    """out_nhwc [b, H, W, cout] = relu(conv3x3(x [b, cin, H, W], w) + bias), fp32 FMA."""
    b, cin, H, W = x.shape
    check(lib().ddpb_conv3x3_small_cin_fwd(ptr(x), ptr(w), ptr(bias), ptr(out_nhwc), cin, w.shape[0], b, H, W,
                                           stream_ptr(stream)), "conv3x3_small_cin_fwd")


def conv3x3_small_cin_wgrad(dact_nhwc, act_nhwc, x, cout, stream=None):
    """(dw [cout, cin, 3, 3], db [cout]) of the convolution above; ReLU' from ``act_nhwc`` (or None)."""
    b, cin, H, W = x.shape
    parts = int(lib().ddpb_conv3x3_small_cin_wgrad_parts(b, H))
    n = cout * 28
    partial = torch.empty(parts * n, dtype=torch.float32, device=x.device)
    out = torch.empty(n, dtype=torch.float32, device=x.device)
    check(lib().ddpb_conv3x3_small_cin_wgrad(ptr(dact_nhwc), ptr(act_nhwc), ptr(x), ptr(partial), ptr(out), cin,
                                             cout, b, H, W, stream_ptr(stream)), "conv3x3_small_cin_wgrad")
    out = out.view(cout, 28)
    return out[:, :27].reshape(cout, cin, 3, 3), out[:, 27].contiguous()


def relu_maxpool2_nhwc_fwd(x_nhwc, bias, out, code, stream=None):
    b, h, w, c = x_nhwc.shape
    check(lib().ddpb_relu_maxpool2_nhwc_fwd(ptr(x_nhwc), ptr(bias), ptr(out), ptr(code), b, c, h, w,
                                            stream_ptr(stream)), "relu_maxpool2_nhwc_fwd")


def relu_maxpool2_nhwc_bwd(dout, code, din_nhwc, stream=None):
    b, h, w, c = din_nhwc.shape
    check(lib().ddpb_relu_maxpool2_nhwc_bwd(ptr(dout), ptr(code), ptr(din_nhwc), b, c, h, w, stream_ptr(stream)),
          "relu_maxpool2_nhwc_bwd")


# ----------------------------------------------------------------------------- small-output linear layers (fp32)
def small_linear_max_out() -> int:
    return int(lib().ddpb_small_linear_max_out())


def small_linear_fwd(x, w, bias, y, stream=None):
    batch, k = x.shape
    check(lib().ddpb_small_linear_fwd(ptr(x), ptr(w), ptr(bias), ptr(y), batch, k, w.shape[0], stream_ptr(stream)),
          "small_linear_fwd")


def small_linear_bwd(dy, x, w, dx, dw, db, stream=None):
    batch, k = x.shape
    nout = w.shape[0]
    scr = torch.empty(int(lib().ddpb_small_linear_bwd_scratch_floats(batch, k, nout)), dtype=torch.float32,
                      device=x.device)
    check(lib().ddpb_small_linear_bwd(ptr(dy), ptr(x), ptr(w), ptr(dx), ptr(dw), ptr(db), ptr(scr), batch, k, nout,
                                      stream_ptr(stream)), "small_linear_bwd")
