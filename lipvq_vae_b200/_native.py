"""ctypes binding of the C-ABI library (include/lipvq_b200.h).

The library is built in-tree (``lipvq_vae_b200/_lib/liblipvq_b200.so``) by ``build_native()``
-- plain ``nvcc -shared`` for sm_100a, no torch types in any signature.  There is NO fallback:
if the library is missing or a call fails, a ``LipVQNativeError`` is raised.
"""
from __future__ import annotations

import ctypes as C
import os
import shutil
import subprocess
import threading

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "_lib")
LIB_PATH = os.environ.get("LIPVQ_LIB") or os.path.join(LIB_DIR, "liblipvq_b200.so")   # LIPVQ_LIB: A/B builds
HEADER = os.path.join(ROOT, "include", "lipvq_b200.h")

SEARCH_EXACT = 0
SEARCH_TENSOR = 1

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-shared", "-Xcompiler", "-fPIC",
]


class LipVQNativeError(RuntimeError):
    pass


class Dims(C.Structure):
    _fields_ = [("F", C.c_int32), ("H1", C.c_int32), ("H", C.c_int32), ("D", C.c_int32), ("K", C.c_int32)]


_PARAM_FIELDS = ["enc0_w", "enc0_b", "enc2_w", "enc2_b", "lat_W", "lat_b", "lat_ci", "codebook",
                 "dec0_w", "dec0_b", "dec2_w", "dec2_b", "out_w", "out_b"]
# reference state_dict name of each struct field (v5:52-68)
PARAM_KEYS = ["encoder.0.weight", "encoder.0.bias", "encoder.2.weight", "encoder.2.bias",
              "to_latent.W", "to_latent.b", "to_latent.ci", "quantizer.codebook",
              "decoder.0.weight", "decoder.0.bias", "decoder.2.weight", "decoder.2.bias",
              "to_output.weight", "to_output.bias"]


class Params(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in _PARAM_FIELDS]


class Grads(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in _PARAM_FIELDS]


_ACT_FIELDS = ["a1", "h1", "a2", "h2", "scale", "Wn", "z_e", "idx", "z_q", "c1", "d1", "c2", "d2", "x_recon", "xr_code", "g_pre"]


class Acts(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in _ACT_FIELDS]


def _sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh"))) + [HEADER]


def needs_build() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(s) > t for s in _sources())


def build_native(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/lipvq_abi.cu for sm_100a into the in-tree shared library."""
    if not force and not needs_build():
        return LIB_PATH
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise LipVQNativeError("nvcc not found; cannot build " + LIB_PATH)
    os.makedirs(LIB_DIR, exist_ok=True)
    tmp = LIB_PATH + ".tmp.%d" % os.getpid()
    extra = os.environ.get("LIPVQ_NVCC_EXTRA", "").split()          # experiment knobs, e.g. -DENC_CH_=64
    cmd = ([nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else [])
           + ["-o", tmp, os.path.join(CSRC, "lipvq_abi.cu")])
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise LipVQNativeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    os.replace(tmp, LIB_PATH)
    return LIB_PATH


_lib = None
_lock = threading.Lock()

_VP, _I64, _I32, _F32, _SZ = C.c_void_p, C.c_int64, C.c_int32, C.c_float, C.c_size_t
_DP, _PP, _GP, _AP = C.POINTER(Dims), C.POINTER(Params), C.POINTER(Grads), C.POINTER(Acts)

# name -> (restype, argtypes); mirrors include/lipvq_b200.h one to one
SIGNATURES = {
    "lipvq_abi_version": (C.c_int, []),
    "lipvq_strerror": (C.c_char_p, [C.c_int]),
    "lipvq_last_cuda_error": (C.c_int, []),
    "lipvq_last_cuda_error_string": (C.c_char_p, []),
    "lipvq_check_device": (C.c_int, []),
    "lipvq_launch_count": (C.c_int64, []),
    "lipvq_profile_begin": (C.c_int, [_I32]),
    "lipvq_profile_end": (C.c_int, [C.c_char_p, _SZ]),
    "lipvq_workspace_bytes": (C.c_int, [_DP, _I64, C.POINTER(_SZ)]),
    "lipvq_encoder_fwd": (C.c_int, [_DP, _PP, _VP, _I64, _AP, _VP, _SZ, _VP]),
    "lipvq_linear_fwd": (C.c_int, [_VP, _VP, _VP, _VP, _I64, _I32, _I32, _I32, _VP]),
    "lipvq_lipschitz_fwd": (C.c_int, [_VP, _VP, _VP, _VP, _VP, _VP, _VP, _I64, _I32, _I32, _VP]),
    "lipvq_vq_argmin": (C.c_int, [_VP, _VP, _I64, _I32, _I32, _I32, _VP, _VP, _VP, _VP, _SZ, _VP]),
    "lipvq_encode_search_fwd": (C.c_int, [_DP, _PP, _VP, _I64, _AP, _I32, _VP, _VP, _SZ, _VP]),
    "lipvq_gather_loss": (C.c_int, [_VP, _VP, _VP, _I64, _I32, _I32, _VP, _VP, _VP, _SZ, _VP]),
    "lipvq_decoder_fwd": (C.c_int, [_DP, _PP, _VP, _VP, _I64, _AP, _VP, _VP, _SZ, _VP]),
    "lipvq_quantize_decode_fwd": (C.c_int, [_DP, _PP, _VP, _I64, _I64, _AP, _VP, _VP, _VP, _SZ, _VP]),
    "lipvq_loss_finalize": (C.c_int, [_VP, _VP, _I64, _I32, _I32, _VP, _VP]),
    "lipvq_decoder_bwd": (C.c_int, [_DP, _PP, _VP, _I64, _I64, _VP, _AP, _GP, _VP, _SZ, _VP]),
    "lipvq_scatter_add": (C.c_int, [_VP, _VP, _I64, _I32, _I32, _VP, _VP, _VP]),
    "lipvq_encoder_bwd": (C.c_int, [_DP, _PP, _VP, _I64, _I64, _VP, _AP, _GP, _VP, _SZ, _VP]),
    "lipvq_forward_fused": (C.c_int, [_DP, _PP, _VP, _I64, _I64, _I32, _VP, _VP, _VP, _VP, _VP, _VP, _SZ, _VP, _VP, _VP, _VP]),
    "lipvq_segment_sum_ws_bytes": (_SZ, [_I64, _I32, _I32]),
    "lipvq_segment_sum": (C.c_int, [_VP, _VP, _VP, _F32, _VP, _I64, _I32, _I32, _I32, _VP, _VP, _VP, _SZ, _VP]),
    "lipvq_forward_small": (C.c_int, [_DP, _PP, _VP, _I64, _I64, _VP, _VP, _VP, _VP, _VP, _VP, _SZ, _VP]),
    "lipvq_forward_small_static": (C.c_int, [_DP, _PP, _VP, _I64, _I64, _VP, _VP, _VP, _VP, _VP, _VP]),
    "lipvq_adamw_step": (C.c_int, [_VP, _VP, _VP, _VP, _I64, _I64, _F32, _F32, _F32, _F32, _F32, _VP]),
    "lipvq_ema_update": (C.c_int, [_VP, _VP, _VP, _VP, _VP, _I32, _I32, _F32, _F32, _VP, _SZ, _VP]),
    "lipvq_linear_bwd_ws_bytes": (_SZ, [_I64, _I32, _I32]),
    "lipvq_linear_bwd": (C.c_int, [_VP, _VP, _VP, _VP, _VP, _VP, _VP, _I64, _I32, _I32, _VP, _SZ, _VP]),
    "lipvq_scatter_add_diff": (C.c_int, [_VP, _VP, _VP, _F32, _VP, _I64, _I32, _I32, _VP, _VP]),
    "lipvq_sse": (C.c_int, [_VP, _VP, _I64, _VP, _VP, _SZ, _VP]),
    "lipvq_vqvae_loss": (C.c_int, [_VP, _VP, _I64, _I32, _I32, _F32, _VP, _VP]),
    "lipvq_vqvae_recon_grad": (C.c_int, [_VP, _VP, _I64, _F32, _VP, _VP, _VP]),
    "lipvq_vqvae_ste_grad": (C.c_int, [_VP, _VP, _VP, _I64, _F32, _VP, _VP]),
}


def lib():
    """Load (once) and return the ctypes handle; raises if the library is absent."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise LipVQNativeError(
                        "native library %s not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                        "there is no CPU fallback" % LIB_PATH)
                h = C.CDLL(LIB_PATH)
                for name, (res, args) in SIGNATURES.items():
                    fn = getattr(h, name)
                    fn.restype = res
                    fn.argtypes = args
                if h.lipvq_abi_version() != 6:
                    raise LipVQNativeError("ABI version mismatch")
                _lib = h
    return _lib


def check(code: int, what: str = "") -> None:
    if code != 0:
        h = lib()
        msg = h.lipvq_strerror(code).decode()
        if code == -5:
            msg += ": " + h.lipvq_last_cuda_error_string().decode()
        raise LipVQNativeError("%s failed (%d): %s" % (what or "lipvq call", code, msg))


def profile_kernels(fn, iters):
    """Run ``fn(i)`` ``iters`` times with the library's per-kernel CUDA-event timing on; returns
    ``{kernel name: (launches, total_ms)}``."""
    h = lib()
    check(h.lipvq_profile_begin(64 * max(iters, 1) + 64), "lipvq_profile_begin")
    try:
        for i in range(iters):
            fn(i)
    finally:
        buf = C.create_string_buffer(1 << 14)
        check(h.lipvq_profile_end(buf, len(buf)), "lipvq_profile_end")
    out = {}
    for line in buf.value.decode().splitlines():
        name, count, ms = line.rsplit(" ", 2)
        out[name] = (int(count), float(ms))
    return out
