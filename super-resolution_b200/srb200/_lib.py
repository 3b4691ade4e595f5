"""ctypes binding of libsrb200.so (include/srb200.h).  No torch types cross this boundary:
tensors are passed as raw device pointers, the stream as a void*.

The library is built in-tree by ``build()`` (nvcc, sm_100a only).  There is no CPU or
torch fallback: if the shared object is missing or a call fails, a RuntimeError whose
text contains "CUDA" is raised (the reference's handlers at esrgan.py:829-836 and
hyper_search.py:108-110 key on that).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import threading

PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(PKG_DIR, "csrc")
# SRB_LIB_PATH: load another build of the same ABI (A/B timing of kernel changes on one GPU box); never rebuilt here
LIB_PATH = os.environ.get("SRB_LIB_PATH") or os.path.join(PKG_DIR, "libsrb200.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]

ACT_NONE, ACT_LRELU, ACT_LRELU_MASK = 0, 1, 2
STORE_PLAIN, STORE_SHUFFLE = 0, 1
IMPL_TCGEN05, IMPL_NAIVE = 0, 1


class Conv3x3Args(C.Structure):
    _fields_ = [
        ("batch", C.c_int), ("H", C.c_int), ("W", C.c_int),
        ("x", C.c_void_p), ("x_cstride", C.c_int), ("x_coff", C.c_int), ("cin", C.c_int),
        ("w", C.c_void_p), ("bias", C.c_void_p), ("cout", C.c_int),
        ("y", C.c_void_p), ("y_cstride", C.c_int), ("y_coff", C.c_int),
        ("store_mode", C.c_int), ("alpha", C.c_float), ("act", C.c_int), ("slope", C.c_float),
        ("mask", C.c_void_p), ("mask_cstride", C.c_int), ("mask_coff", C.c_int),
        ("res1", C.c_void_p), ("res1_cstride", C.c_int), ("res1_coff", C.c_int), ("res1_scale", C.c_float),
        ("res2", C.c_void_p), ("res2_cstride", C.c_int), ("res2_coff", C.c_int), ("res2_scale", C.c_float),
        ("impl", C.c_int), ("flags", C.c_int),
        ("sign_out", C.c_void_p), ("sign_cstride", C.c_int), ("sign_coff", C.c_int),
        ("mask_bits", C.c_void_p), ("maskb_cstride", C.c_int), ("maskb_coff", C.c_int),
        ("colsum", C.c_void_p),
        ("res_f32", C.c_int), ("y32", C.c_void_p), ("y32_cstride", C.c_int), ("y32_coff", C.c_int),
        ("res3", C.c_void_p), ("res3_cstride", C.c_int), ("res3_coff", C.c_int), ("res3_scale", C.c_float),
        ("chan_scale", C.c_void_p),
    ]


class Wgrad3x3Args(C.Structure):
    _fields_ = [
        ("batch", C.c_int), ("H", C.c_int), ("W", C.c_int),
        ("x", C.c_void_p), ("x_cstride", C.c_int), ("x_coff", C.c_int), ("cin", C.c_int),
        ("gy", C.c_void_p), ("gy_cstride", C.c_int), ("gy_coff", C.c_int), ("cout", C.c_int),
        ("dw", C.c_void_p), ("dbias", C.c_void_p), ("alpha", C.c_float),
        ("accumulate", C.c_int), ("unshuffle", C.c_int),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t),
        ("impl", C.c_int), ("flags", C.c_int), ("gy_colsum", C.c_void_p),
    ]


class PackItem(C.Structure):
    _fields_ = [("w", C.c_void_p), ("wp", C.c_void_p), ("cout", C.c_int), ("cin", C.c_int), ("ci_off", C.c_int),
                ("n_out", C.c_int), ("flag", C.c_int), ("kind", C.c_int), ("kb_dst", C.c_int), ("kb_total", C.c_int),
                ("scale", C.c_float), ("pad", C.c_int)]


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh")))


def needs_build() -> bool:
    if os.environ.get("SRB_LIB_PATH"):
        return False
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    hdr = os.path.join(os.path.dirname(PKG_DIR), "include", "srb200.h")
    return any(os.path.getmtime(s) > t for s in sources() + [hdr])


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/srb200_all.cu (unity build) into libsrb200.so for sm_100a."""
    if not force and not needs_build():
        return LIB_PATH
    nvcc = os.environ.get("NVCC", "nvcc")
    tmp = LIB_PATH + ".tmp%d" % os.getpid()          # never leave a half-written library where lib() would load it
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
          (["-DSRB_EXPERIMENTAL"] if os.environ.get("SRB_EXPERIMENTAL") == "1" else []) + \
          ["-o", tmp, os.path.join(CSRC, "srb200_all.cu")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        if os.path.exists(tmp):
            os.remove(tmp)
        raise RuntimeError("CUDA extension build failed:\n" + r.stdout + r.stderr)
    os.replace(tmp, LIB_PATH)
    if verbose:
        print(r.stderr)
    return LIB_PATH


_lib = None
_lock = threading.Lock()

_SIGS = {
    "srb_last_error": (C.c_char_p, []),
    "srb_abi_version": (C.c_int, []),
    "srb_has_experimental": (C.c_int, []),
    "srb_read_fault": (C.c_int, [C.POINTER(C.c_uint)]),
    "srb_conv3x3": (C.c_int, [C.POINTER(Conv3x3Args), C.c_void_p]),
    "srb_conv3x3_chain": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]),
    "srb_conv3x3_chain_flag_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int, C.c_int]),
    "srb_pack_conv3x3_fwd": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p]),
    "srb_pack_conv3x3_dgrad": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float,
                                         C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "srb_pack_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "srb_wgrad3x3": (C.c_int, [C.POINTER(Wgrad3x3Args), C.c_void_p]),
    "srb_wgrad3x3_group": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "srb_wgrad3x3_static": (C.c_int, [C.POINTER(Wgrad3x3Args), C.c_void_p]),
    "srb_wgrad3x3_static_max_bytes": (C.c_size_t, []),
    "srb_wgrad3x3_group_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "srb_wgrad3x3_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "srb_conv_in_fwd": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p,
                                  C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int,
                                  C.c_void_p]),
    "srb_conv_in_bwd": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p,
                                  C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                  C.c_void_p]),
    "srb_conv_out_fwd": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                   C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_int, C.c_void_p,
                                   C.c_void_p, C.c_void_p, C.c_void_p]),
    "srb_conv_out_bwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float,
                                   C.c_float, C.c_float, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int,
                                   C.c_float, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                   C.c_void_p]),
    "srb_unshuffle2": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                 C.c_int, C.c_int, C.c_void_p]),
    "srb_chan_mask_split": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int,
                                      C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "srb_dconv_fwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                C.c_int, C.c_int, C.c_int, C.c_float, C.c_void_p]),
    "srb_set_dconv_impl": (C.c_int, [C.c_int]),
    "srb_dconv_dgrad": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                  C.c_int, C.c_void_p]),
    "srb_dconv_wgrad": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                  C.c_int, C.c_int, C.c_void_p]),
    "srb_lrelu_fwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_longlong, C.c_float, C.c_void_p]),
    "srb_lrelu_bwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong, C.c_float, C.c_void_p]),
    "srb_sumpool_fwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "srb_sumpool_bwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "srb_dmma_conv": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                C.c_float, C.c_void_p]),
    "srb_dmma_wgrad": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                 C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "srb_dl1_fwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                              C.c_int, C.c_int, C.c_float, C.c_void_p]),
    "srb_dl1_dgrad": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "srb_dl1_wgrad": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                C.c_void_p]),
    "srb_dpack": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                            C.c_void_p, C.c_void_p]),
    "srb_adam_step": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong, C.c_float, C.c_float,
                                C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_float,
                                C.c_void_p]),
    "srb_pixel_losses": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float,
                                   C.c_float, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "srb_pixel_losses_pow": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float,
                                       C.c_float, C.c_float, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_void_p]),
    "srb_soft_hist": (C.c_int, [C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_void_p, C.c_void_p]),
    "srb_kld_hist": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "srb_soft_hist_bwd": (C.c_int, [C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_void_p,
                                    C.c_float, C.c_int, C.c_void_p, C.c_void_p]),
    "srb_nnz_loss": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, C.c_void_p,
                               C.c_void_p, C.c_void_p, C.c_void_p]),
    "srb_mask_loss": (C.c_int, [C.c_void_p, C.c_void_p, C.c_longlong, C.c_float, C.c_float, C.c_int, C.c_void_p,
                                C.c_void_p, C.c_void_p]),
    "srb_hit_loss": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float,
                               C.c_float, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "srb_sinkhorn_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "srb_sinkhorn_loss": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_int, C.c_float, C.c_int,
                                    C.c_int, C.c_float, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p]),
    "srb_ragan": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p,
                            C.c_void_p]),
    "srb_gp_fwd": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]),
    "srb_gp_bwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]),
    "srb_extract_batch": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_void_p]),
    "srb_image_stats": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_int, C.c_void_p, C.c_void_p]),
    "srb_l1_metrics": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                 C.c_void_p]),
    "srb_eval_rasters": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int,
                                   C.c_void_p, C.c_void_p, C.c_void_p]),
    "srb_igemm_read_stamps": (C.c_int, [C.c_void_p]),
    "srb_probe_mma_rate": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                     C.c_void_p]),
}


def exported_symbols():
    return list(_SIGS)


def lib():
    """Load libsrb200.so (once).  Fails loudly when it is absent: there is no fallback path."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise RuntimeError(
                        "CUDA extension libsrb200.so is not built (run `python -c 'import __graft_entry__ as g; "
                        "g.build()'`); super-resolution_b200 has no CPU fallback")
                l = C.CDLL(LIB_PATH)
                for name, (res, args) in _SIGS.items():
                    fn = getattr(l, name)
                    fn.restype = res
                    fn.argtypes = args
                if l.srb_abi_version() != 3:
                    raise RuntimeError("CUDA extension libsrb200.so is stale (ABI %d, bindings expect 3): rebuild it"
                                       % l.srb_abi_version())
                _lib = l
    return _lib


CALLS = [0]       # successful C-ABI launches (every entry point launches >= 1 kernel); bench.py reports the delta


def check(status: int, what: str = "") -> None:
    CALLS[0] += 1
    if status != 0:
        msg = lib().srb_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{msg} [{what}]" if what else msg)
