"""ctypes binding of libcss_b200.so (the C-ABI declared in include/css_b200.h).

There is NO fallback: if the library is missing or CUDA is unavailable the product path raises.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CSS_B200_LIB") or os.path.join(_HERE, "csrc", "libcss_b200.so")   # override: A/B of kernel builds
KP = 264
NBINS = 257

_lib = None


class NativeError(RuntimeError):
    pass


def lib():
    """Load (once) and return the shared library.  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NativeError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(css_b200 has no CPU or PyTorch fallback)"
            )
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.css_last_error.restype = ctypes.c_char_p
        _lib.css_launch_count.restype = ctypes.c_ulonglong
        _lib.css_conformer_create.restype = ctypes.c_void_p
        _lib.css_conformer_workspace_bytes.restype = ctypes.c_size_t
        _lib.css_blstm_create.restype = ctypes.c_void_p
        _lib.css_blstm_workspace_bytes.restype = ctypes.c_size_t
        _lib.css_gemm_tc_scratch_bytes.restype = ctypes.c_size_t
        if os.environ.get("CSS_B200_GEMM_MODE"):
            if _lib.css_set_gemm_mode(int(os.environ["CSS_B200_GEMM_MODE"])) != 0:
                raise NativeError(_lib.css_last_error().decode())
    return _lib


def require_cuda(t, name="tensor"):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise NativeError(f"{name} must be a CUDA tensor: css_b200 runs only on the GPU (no CPU fallback)")
    return t


class DevPtr(ctypes.c_void_p):
    """A device pointer that remembers which GPU it lives on, so that ``call`` can launch on that GPU."""

    device = None


def ptr(t):
    p = DevPtr(0 if t is None else t.data_ptr())
    if t is not None and t.is_cuda:
        p.device = t.device.index
    return p


def stream(device=None):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def check(rc, what=""):
    if rc != 0:
        msg = lib().css_last_error().decode("utf-8", "replace")
        raise NativeError(f"{what} failed ({rc}): {msg}")


def call(name, *args):
    """Call an `int css_xxx(...)` entry point, appending the current stream, raising on error.

    The kernels are launched on the device that owns the tensors (all pointer arguments must live on one GPU), on that
    device's current stream -- not on whatever device happens to be current in the calling thread."""
    fn = getattr(lib(), name)
    devs = {a.device for a in args if isinstance(a, DevPtr) and a.device is not None}
    if len(devs) > 1:
        raise NativeError(f"{name}: tensors live on different GPUs {sorted(devs)}")
    dev = devs.pop() if devs else torch.cuda.current_device()
    if dev == torch.cuda.current_device():
        check(fn(*args, stream()), name)
    else:
        with torch.cuda.device(dev):
            check(fn(*args, stream(dev)), name)


def set_gemm_mode(mode):
    """0 = fp32 SIMT, 1 = tcgen05 bf16x3 (fp32-class accuracy), 2 = tcgen05 bf16."""
    check(lib().css_set_gemm_mode(int(mode)), "css_set_gemm_mode")


def get_gemm_mode():
    return int(lib().css_get_gemm_mode())


def launch_count(reset=False):
    return int(lib().css_launch_count(1 if reset else 0))


c_int = ctypes.c_int
c_long = ctypes.c_long
c_float = ctypes.c_float
c_size_t = ctypes.c_size_t


# ---- descriptor structs mirrored from include/css_b200.h ----
class FfnWeights(ctypes.Structure):
    _fields_ = [(n, ctypes.c_void_p) for n in ("ln_g", "ln_b", "w1", "b1", "w2", "b2")]


class ConformerBlock(ctypes.Structure):
    _fields_ = (
        [("ff_in", FfnWeights)]
        + [(n, ctypes.c_void_p) for n in ("att_ln_g", "att_ln_b", "wqkv", "bqkv", "wo", "bo")]
        + [(n, ctypes.c_void_p) for n in ("conv_ln_g", "conv_ln_b", "dw_w", "dw_b", "bn_scale", "bn_shift")]
        + [(n, ctypes.c_float) for n in ("pw1_w0", "pw1_b0", "pw1_w1", "pw1_b1", "pw2_w", "pw2_b")]
        + [("ff_out", FfnWeights)]
        + [(n, ctypes.c_void_p) for n in ("out_ln_g", "out_ln_b")]
    )


class ConformerDesc(ctypes.Structure):
    _fields_ = (
        [(n, ctypes.c_int) for n in ("idim", "ldin", "d_model", "heads", "d_ffn", "num_blocks", "num_bins", "num_out", "pe_rows")]
        + [(n, ctypes.c_void_p) for n in ("embed_w", "embed_b", "embed_ln_g", "embed_ln_b", "pe", "head_w", "head_b")]
    )


class BlstmLayer(ctypes.Structure):
    _fields_ = [(n, ctypes.c_void_p) for n in ("w_ih", "b_gates", "w_hh_f", "w_hh_b", "ln_g", "ln_b")]


class BlstmDesc(ctypes.Structure):
    _fields_ = (
        [(n, ctypes.c_int) for n in ("idim", "ldin", "hdim", "num_layers", "num_bins", "num_out")]
        + [(n, ctypes.c_void_p) for n in ("embed_w", "embed_b", "embed_ln_g", "embed_ln_b", "head_w", "head_b")]
    )
