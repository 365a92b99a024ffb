"""ctypes binding of libmsda_b200.so (include/msda_b200.h).

The library is the product; there is no fallback.  If it is missing, or a call
fails, this module raises -- it never routes around the CUDA kernels.
"""
import contextlib
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmsda_b200.so")

MSDA_F32, MSDA_BF16, MSDA_F64, MSDA_F16 = 0, 1, 2, 3
_DTYPE_CODE = {torch.float32: MSDA_F32, torch.bfloat16: MSDA_BF16,
               torch.float64: MSDA_F64, torch.float16: MSDA_F16}


class MsdaDims(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in ("N", "S", "M", "D", "L", "Lq", "P")]


class MsdaFusedInputs(ctypes.Structure):
    _fields_ = [("sampling_offsets", ctypes.c_void_p), ("attn_logits", ctypes.c_void_p),
                ("reference_points", ctypes.c_void_p), ("offset_norm", ctypes.c_void_p),
                ("query_mask", ctypes.c_void_p), ("ref_dim", ctypes.c_int32)]


MSDA_E_SHAPE = -2
MSDA_E_SCRATCH_LIMIT = -6
STATIC_SCRATCH_BYTES = 32 * 1024 * 1024       # MSDA_B200_STATIC_SCRATCH_BYTES

EXPORTS = ("msda_b200_abi_version", "msda_b200_error_string", "msda_b200_forward",
           "msda_b200_forward_fused", "msda_b200_backward_fused",
           "msda_b200_backward_workspace_bytes", "msda_b200_backward", "msda_b200_backward_static_scratch",
           "msda_b200_last_launch_count", "msda_b200_trace_next")
# include/kinet_layers_b200.h (glue kernels of the layers that call the op)
LAYER_EXPORTS = ("kinet_b200_add_layernorm_forward", "kinet_b200_add_layernorm_backward",
                 "kinet_b200_colsum_partial_floats", "kinet_b200_colsum", "kinet_b200_relu_backward_colsum",
                 "kinet_b200_small_mha_forward", "kinet_b200_small_mha_backward",
                 "kinet_b200_dropout_forward", "kinet_b200_relu_dropout_backward_colsum")
MHA_MAX_D = 64                                 # KINET_MHA_MAX_D


class KinetLnFwdArgs(ctypes.Structure):
    _fields_ = [("x", ctypes.c_void_p), ("x_dtype", ctypes.c_int32),
                ("r", ctypes.c_void_p), ("r_dtype", ctypes.c_int32),
                ("pos", ctypes.c_void_p), ("pos_dtype", ctypes.c_int32),
                ("pos_rows", ctypes.c_int64),
                ("gamma", ctypes.c_void_p), ("beta", ctypes.c_void_p),
                ("y", ctypes.c_void_p), ("y_dtype", ctypes.c_int32),
                ("y_lp", ctypes.c_void_p),
                ("ypos_lp", ctypes.c_void_p), ("lp_dtype", ctypes.c_int32),
                ("mean_rstd", ctypes.c_void_p),
                ("rows", ctypes.c_int64), ("C", ctypes.c_int32),
                ("eps", ctypes.c_float),
                ("drop_p", ctypes.c_float), ("drop_site", ctypes.c_uint32),
                ("drop_seed", ctypes.c_uint64), ("drop_step", ctypes.c_void_p)]


class KinetLnBwdArgs(ctypes.Structure):
    _fields_ = [("dy", ctypes.c_void_p), ("dy_dtype", ctypes.c_int32),
                ("dy_lp", ctypes.c_void_p), ("dypos_lp", ctypes.c_void_p), ("lp_dtype", ctypes.c_int32),
                ("x", ctypes.c_void_p), ("x_dtype", ctypes.c_int32),
                ("r", ctypes.c_void_p), ("r_dtype", ctypes.c_int32),
                ("gamma", ctypes.c_void_p),
                ("mean_rstd", ctypes.c_void_p),
                ("dx", ctypes.c_void_p),
                ("dr", ctypes.c_void_p),
                ("dgamma_dbeta", ctypes.c_void_p),
                ("partial", ctypes.c_void_p),
                ("rows", ctypes.c_int64), ("C", ctypes.c_int32),
                ("want_dbias", ctypes.c_int32),
                ("drop_p", ctypes.c_float), ("drop_site", ctypes.c_uint32),
                ("drop_seed", ctypes.c_uint64), ("drop_step", ctypes.c_void_p)]



class KinetMhaArgs(ctypes.Structure):
    _fields_ = [("q", ctypes.c_void_p), ("k", ctypes.c_void_p), ("v", ctypes.c_void_p),
                ("q_batch_stride", ctypes.c_int64), ("q_row_stride", ctypes.c_int64),
                ("k_batch_stride", ctypes.c_int64), ("k_row_stride", ctypes.c_int64),
                ("v_batch_stride", ctypes.c_int64), ("v_row_stride", ctypes.c_int64),
                ("key_padding_mask", ctypes.c_void_p),
                ("out", ctypes.c_void_p),
                ("out_batch_stride", ctypes.c_int64), ("out_row_stride", ctypes.c_int64),
                ("lse", ctypes.c_void_p),
                ("dtype", ctypes.c_int32),
                ("N", ctypes.c_int32), ("H", ctypes.c_int32), ("D", ctypes.c_int32),
                ("Lq", ctypes.c_int32), ("Lk", ctypes.c_int32),
                ("scale", ctypes.c_float)]


class KinetMhaBwdArgs(ctypes.Structure):
    _fields_ = [("fwd", KinetMhaArgs),
                ("dout", ctypes.c_void_p),
                ("dout_batch_stride", ctypes.c_int64), ("dout_row_stride", ctypes.c_int64),
                ("dq", ctypes.c_void_p), ("dk", ctypes.c_void_p), ("dv", ctypes.c_void_p)]


_lib = None


def load():
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "kinet_b200: %s is missing -- build it with `python -m kinet_b200.build` "
            "(there is no CPU or PyTorch fallback for MSDeformAttn)" % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    vp, i, dp = ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(MsdaDims)
    lib.msda_b200_abi_version.restype = i
    lib.msda_b200_error_string.restype = ctypes.c_char_p
    lib.msda_b200_error_string.argtypes = [i]
    lib.msda_b200_forward.restype = i
    lib.msda_b200_forward.argtypes = [vp, i, vp, vp, vp, vp, i, vp, dp, vp]
    lib.msda_b200_backward_workspace_bytes.restype = ctypes.c_size_t
    lib.msda_b200_backward_workspace_bytes.argtypes = [i, dp]
    lib.msda_b200_backward.restype = i
    lib.msda_b200_backward.argtypes = [vp, i, vp, vp, vp, vp, i, vp, vp, vp, vp, vp,
                                       ctypes.c_size_t, dp, vp]
    lib.msda_b200_backward_static_scratch.restype = i
    lib.msda_b200_backward_static_scratch.argtypes = [vp, i, vp, vp, vp, vp, i, vp, vp, vp, vp, dp, vp]
    fp = ctypes.POINTER(MsdaFusedInputs)
    lib.msda_b200_forward_fused.restype = i
    lib.msda_b200_forward_fused.argtypes = [vp, i, vp, vp, fp, i, vp, dp, vp]
    lib.msda_b200_backward_fused.restype = i
    lib.msda_b200_backward_fused.argtypes = [vp, i, vp, vp, fp, i, vp, vp, vp, vp, vp, vp, ctypes.c_size_t, dp, vp]
    lib.msda_b200_last_launch_count.restype = i
    lib.msda_b200_trace_next.restype = None
    lib.msda_b200_trace_next.argtypes = [vp, vp]
    lib.kinet_b200_add_layernorm_forward.restype = i
    lib.kinet_b200_add_layernorm_forward.argtypes = [ctypes.POINTER(KinetLnFwdArgs), vp]
    lib.kinet_b200_add_layernorm_backward.restype = i
    lib.kinet_b200_add_layernorm_backward.argtypes = [ctypes.POINTER(KinetLnBwdArgs), vp]
    lib.kinet_b200_colsum_partial_floats.restype = ctypes.c_size_t
    lib.kinet_b200_colsum_partial_floats.argtypes = [ctypes.c_int32]
    lib.kinet_b200_colsum.restype = i
    lib.kinet_b200_colsum.argtypes = [vp, ctypes.c_int32, ctypes.c_int64, ctypes.c_int32, vp, ctypes.c_int32, vp, vp]
    lib.kinet_b200_relu_backward_colsum.restype = i
    lib.kinet_b200_relu_backward_colsum.argtypes = [vp, vp, ctypes.c_int32, ctypes.c_int64, ctypes.c_int32, vp, vp,
                                                    ctypes.c_int32, vp, vp]
    lib.kinet_b200_dropout_forward.restype = i
    lib.kinet_b200_dropout_forward.argtypes = [vp, vp, ctypes.c_int32, ctypes.c_int64, ctypes.c_float, ctypes.c_uint32,
                                               ctypes.c_uint64, vp, vp]
    lib.kinet_b200_relu_dropout_backward_colsum.restype = i
    lib.kinet_b200_relu_dropout_backward_colsum.argtypes = [vp, vp, ctypes.c_float, ctypes.c_int32, ctypes.c_int64,
                                                            ctypes.c_int32, vp, vp, ctypes.c_int32, vp, vp]
    lib.kinet_b200_small_mha_forward.restype = i
    lib.kinet_b200_small_mha_forward.argtypes = [ctypes.POINTER(KinetMhaArgs), vp]
    lib.kinet_b200_small_mha_backward.restype = i
    lib.kinet_b200_small_mha_backward.argtypes = [ctypes.POINTER(KinetMhaBwdArgs), vp]
    if lib.msda_b200_abi_version() != 3:
        raise RuntimeError("kinet_b200: libmsda_b200.so ABI version mismatch")
    _lib = lib
    return lib


_NO_GUARD = contextlib.nullcontext()


def raw_stream(device):
    """cudaStream_t (as int) of torch's current stream on `device`: one C call, no Stream object."""
    try:
        return torch._C._cuda_getCurrentRawStream(device.index)
    except AttributeError:          # older / newer torch without the private accessor
        return torch.cuda.current_stream(device).cuda_stream


def device_guard(device):
    """Kernels launch on the CUDA runtime's current device: switch only when it is not `device`."""
    return _NO_GUARD if device.index == torch.cuda.current_device() else torch.cuda.device(device)


def dtype_code(dtype):
    try:
        return _DTYPE_CODE[dtype]
    except KeyError:
        raise RuntimeError("kinet_b200: unsupported dtype %s" % dtype)


def check(rc, what):
    if rc != 0:
        msg = load().msda_b200_error_string(rc).decode()
        raise RuntimeError("%s failed: %s (code %d)" % (what, msg, rc))


def last_launch_count():
    return int(load().msda_b200_last_launch_count())
