"""ctypes binding of ``libb200_w4a16.so`` (the C ABI in ``include/b200_w4a16.h``).

There is no fallback: if the library is missing or cannot be loaded every op raises.
``build()`` compiles it in-tree with nvcc for sm_100a (works without a GPU).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Optional

_PKG = os.path.dirname(os.path.abspath(__file__))
# The product library.  Tuning harnesses under scripts/ that want an instrumented build
# (`make -C csrc tune | trace | diag`) call use_library() before the first op; nothing here
# reads the environment.
LIB_PATH = os.path.join(_PKG, "lib", "libb200_w4a16.so")
CSRC = os.path.join(_PKG, "csrc")


_lib: Optional[ctypes.CDLL] = None


def use_library(file_name: str) -> None:
    """Tuning aid (scripts/ only): load lib/<file_name> instead of the product library."""
    global LIB_PATH
    if _lib is not None:
        raise B200LibraryError("use_library() must be called before the library is loaded")
    LIB_PATH = os.path.join(_PKG, "lib", file_name)

B200_OK, B200_ERR_BAD_SHAPE, B200_ERR_BAD_ARG, B200_ERR_WORKSPACE, B200_ERR_CUDA, \
    B200_ERR_NO_DEVICE, B200_ERR_STATIC_LIMIT = range(7)
LAYOUT_AWQ, LAYOUT_B200 = 0, 1
REPACK_AWQ_TO_B200, REPACK_B200_TO_AWQ = 0, 1

# every symbol include/b200_w4a16.h declares
EXPORTS = (
    "b200_w4a16_version", "b200_w4a16_strerror", "b200_w4a16_last_cuda_error",
    "b200_w4a16_repacked_words", "b200_w4a16_workspace_bytes", "b200_w4a16_gemm",
    "b200_w4a16_dequant", "b200_w4a16_repack", "b200_w4a16_gemm_plan", "b200_w4a16_gemm_variant", "b200_w4a16_gemm_unit_tiles", "b200_w4a16_gemm_cta_pair",
    "b200_w4a16_ar_region_bytes", "b200_w4a16_ar_alloc", "b200_w4a16_ar_free",
    "b200_w4a16_ar_export", "b200_w4a16_ar_import", "b200_w4a16_ar_unimport",
    "b200_w4a16_allreduce", "b200_w4a16_allreduce_twoshot", "b200_w4a16_gemm_allreduce", "b200_w4a16_gemm_gated",
    "b200_w4a16_moe_gather", "b200_w4a16_moe_scatter_add", "b200_w4a16_moe_grouped_gemm", "b200_w4a16_moe_combine", "b200_w4a16_moe_route", "b200_w4a16_gemm_noalloc", "b200_w4a16_gemm_rows", "b200_w4a16_moe_compact", "b200_w4a16_moe_gather_rows", "b200_w4a16_moe_scatter_add_rows", "b200_w4a16_silu_and_mul_rows", "b200_w4a16_silu_and_mul", "b200_w4a16_gemm_silu_mul", "b200_w4a16_gemm_silu_mul_gated", "b200_w4a16_add_rms_norm", "b200_w4a16_allreduce_add_rms_norm",
    "b200_w4a16_allreduce_add_rms_norm_ll", "b200_w4a16_allreduce_ll", "b200_w4a16_workspace_init",
)


class B200LibraryError(RuntimeError):
    pass


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile the CUDA library in-tree (``make -C csrc``)."""
    cmd = ["make", "-C", CSRC, "-j4"] + (["-B"] if force else [])
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout)
        print(res.stderr)
    if res.returncode != 0:
        raise B200LibraryError("building libb200_w4a16.so failed (see output above)")
    return LIB_PATH


def lib() -> ctypes.CDLL:
    """Load the library once; raise loudly if it is not there."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200LibraryError(
            f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C vllm-quick_b200/csrc`. This package has no CPU or PyTorch fallback.")
    L = ctypes.CDLL(LIB_PATH)
    vp, i32, sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t
    L.b200_w4a16_version.restype = i32
    L.b200_w4a16_strerror.restype = ctypes.c_char_p
    L.b200_w4a16_strerror.argtypes = [i32]
    L.b200_w4a16_last_cuda_error.restype = i32
    L.b200_w4a16_repacked_words.restype = sz
    L.b200_w4a16_repacked_words.argtypes = [i32, i32]
    L.b200_w4a16_workspace_bytes.restype = sz
    L.b200_w4a16_workspace_bytes.argtypes = [i32, i32, i32, i32]
    L.b200_w4a16_workspace_init.restype = i32
    L.b200_w4a16_workspace_init.argtypes = [vp, sz, vp]
    L.b200_w4a16_gemm.restype = i32
    L.b200_w4a16_gemm.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp, sz, vp]
    L.b200_w4a16_gemm_gated.restype = i32
    L.b200_w4a16_gemm_gated.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp, vp, sz, vp]
    L.b200_w4a16_dequant.restype = i32
    L.b200_w4a16_dequant.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, vp]
    L.b200_w4a16_repack.restype = i32
    L.b200_w4a16_repack.argtypes = [vp, vp, i32, i32, i32, vp]
    L.b200_w4a16_gemm_plan.restype = i32
    L.b200_w4a16_gemm_plan.argtypes = [i32, i32, i32, i32, i32] + [ctypes.POINTER(i32)] * 4
    L.b200_w4a16_ar_region_bytes.restype = sz
    L.b200_w4a16_ar_region_bytes.argtypes = [sz, i32, i32]
    L.b200_w4a16_gemm_allreduce.restype = i32
    L.b200_w4a16_gemm_allreduce.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, sz,
                                            ctypes.POINTER(vp), i32, i32, sz, i32, vp]
    L.b200_w4a16_ar_alloc.restype = i32
    L.b200_w4a16_ar_alloc.argtypes = [sz, ctypes.POINTER(vp)]
    L.b200_w4a16_ar_free.restype = i32
    L.b200_w4a16_ar_free.argtypes = [vp]
    L.b200_w4a16_ar_export.restype = i32
    L.b200_w4a16_ar_export.argtypes = [vp, vp]
    L.b200_w4a16_ar_import.restype = i32
    L.b200_w4a16_ar_import.argtypes = [vp, ctypes.POINTER(vp)]
    L.b200_w4a16_ar_unimport.restype = i32
    L.b200_w4a16_ar_unimport.argtypes = [vp]
    L.b200_w4a16_allreduce.restype = i32
    L.b200_w4a16_allreduce.argtypes = [ctypes.POINTER(vp), i32, i32, vp, vp, sz, sz, i32, vp]
    L.b200_w4a16_allreduce_twoshot.restype = i32
    L.b200_w4a16_allreduce_twoshot.argtypes = [ctypes.POINTER(vp), i32, i32, vp, vp, sz, sz, i32, vp]
    L.b200_w4a16_gemm_variant.restype = i32
    L.b200_w4a16_gemm_variant.argtypes = [i32, i32, i32, i32]
    L.b200_w4a16_gemm_unit_tiles.restype = i32
    L.b200_w4a16_gemm_unit_tiles.argtypes = [i32, i32, i32, i32]
    L.b200_w4a16_gemm_cta_pair.restype = i32
    L.b200_w4a16_gemm_cta_pair.argtypes = [i32, i32, i32, i32]
    f32 = ctypes.c_float
    L.b200_w4a16_moe_gather.restype = i32
    L.b200_w4a16_moe_gather.argtypes = [vp, vp, vp, sz, i32, vp]
    L.b200_w4a16_moe_scatter_add.restype = i32
    L.b200_w4a16_moe_scatter_add.argtypes = [vp, vp, vp, vp, sz, i32, vp]
    L.b200_w4a16_moe_grouped_gemm.restype = i32
    L.b200_w4a16_moe_grouped_gemm.argtypes = [vp, ctypes.POINTER(vp), ctypes.POINTER(vp), ctypes.POINTER(vp),
                                              vp, i32, i32, i32, i32, i32, vp, sz, i32, vp, sz, vp]
    L.b200_w4a16_moe_route.restype = i32
    L.b200_w4a16_moe_route.argtypes = [vp, vp, vp, vp, vp, vp, sz, i32, i32, i32, vp]
    L.b200_w4a16_gemm_rows.restype = i32
    L.b200_w4a16_gemm_rows.argtypes = [vp, vp, vp, vp, vp, i32, vp, i32, i32, i32, vp, sz, vp]
    L.b200_w4a16_moe_compact.restype = i32
    L.b200_w4a16_moe_compact.argtypes = [vp, vp, vp, vp, sz, i32, i32, i32, vp]
    L.b200_w4a16_moe_gather_rows.restype = i32
    L.b200_w4a16_moe_gather_rows.argtypes = [vp, vp, vp, vp, sz, i32, vp]
    L.b200_w4a16_moe_scatter_add_rows.restype = i32
    L.b200_w4a16_moe_scatter_add_rows.argtypes = [vp, vp, vp, vp, vp, sz, i32, vp]
    L.b200_w4a16_silu_and_mul_rows.restype = i32
    L.b200_w4a16_silu_and_mul_rows.argtypes = [vp, vp, sz, vp, i32, i32, vp]
    L.b200_w4a16_gemm_noalloc.restype = i32
    L.b200_w4a16_gemm_noalloc.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, vp]
    L.b200_w4a16_moe_combine.restype = i32
    L.b200_w4a16_moe_combine.argtypes = [vp, vp, vp, vp, i32, i32, sz, i32, vp]
    L.b200_w4a16_silu_and_mul.restype = i32
    L.b200_w4a16_silu_and_mul.argtypes = [vp, vp, sz, i32, i32, vp]
    L.b200_w4a16_gemm_silu_mul.restype = i32
    L.b200_w4a16_gemm_silu_mul.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, sz, vp]
    L.b200_w4a16_gemm_silu_mul_gated.restype = i32
    L.b200_w4a16_gemm_silu_mul_gated.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp, sz, vp]
    L.b200_w4a16_add_rms_norm.restype = i32
    L.b200_w4a16_add_rms_norm.argtypes = [vp, vp, vp, vp, vp, f32, sz, i32, vp]
    L.b200_w4a16_allreduce_add_rms_norm.restype = i32
    L.b200_w4a16_allreduce_add_rms_norm.argtypes = [ctypes.POINTER(vp), i32, i32, vp, vp, vp, vp, vp, f32,
                                              sz, i32, sz, i32, vp]
    L.b200_w4a16_allreduce_add_rms_norm_ll.restype = i32
    L.b200_w4a16_allreduce_add_rms_norm_ll.argtypes = [ctypes.POINTER(vp), i32, i32, vp, vp, vp, vp, vp,
                                                       f32, sz, i32, sz, sz, i32, vp]
    L.b200_w4a16_allreduce_ll.restype = i32
    L.b200_w4a16_allreduce_ll.argtypes = [ctypes.POINTER(vp), i32, i32, vp, vp, sz, sz, sz, i32, vp]
    _lib = L
    return L


def check(status: int, what: str) -> None:
    """Map a b200_status_t to the exception the reference op would raise
    (std::invalid_argument -> ValueError, csrc/quantization/awq/gemm_kernels.cu:412-419)."""
    if status == B200_OK:
        return
    msg = lib().b200_w4a16_strerror(status).decode()
    if status == B200_ERR_BAD_SHAPE:
        raise ValueError(f"{what}: {msg}")
    if status == B200_ERR_CUDA:
        raise RuntimeError(f"{what}: {msg} (cudaError {lib().b200_w4a16_last_cuda_error()})")
    raise RuntimeError(f"{what}: {msg}")


def gemm_plan(M: int, K: int, N: int, G: int = 128, split_k_hint: int = 0) -> dict:
    tok, split, grid, smem = (ctypes.c_int() for _ in range(4))
    check(lib().b200_w4a16_gemm_plan(M, K, N, G, split_k_hint, ctypes.byref(tok),
                                     ctypes.byref(split), ctypes.byref(grid),
                                     ctypes.byref(smem)), "gemm_plan")
    variant = lib().b200_w4a16_gemm_variant(M, K, N, G)
    return {"tok_tile": tok.value, "split_k": split.value, "grid": grid.value,
            "smem_bytes": smem.value,
            "variant": {0: "tcgen05", 1: "decode"}.get(variant, str(variant)),
            "unit_tiles": lib().b200_w4a16_gemm_unit_tiles(M, K, N, G),
            "cta_pair": lib().b200_w4a16_gemm_cta_pair(M, K, N, G) > 0,
            "token_tiles": max(1, lib().b200_w4a16_gemm_cta_pair(M, K, N, G))}
