"""ctypes binding of ``libaar_cuda.so`` (C ABI in ``include/aar_b200.h``).

The CUDA path fails loudly: if the library cannot be built or loaded on a machine that has a
GPU, importing this module raises instead of falling back to eager PyTorch.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

from .. import build as _build

MAX_RANKS = 16
F32, BF16 = 0, 1
ALGO_TWOSHOT, ALGO_ONESHOT, ALGO_NVLS, ALGO_DIRECT = 0, 1, 2, 3
ALGO_NAMES = {"twoshot": ALGO_TWOSHOT, "oneshot": ALGO_ONESHOT, "nvls": ALGO_NVLS, "direct": ALGO_DIRECT}
OP_ALLREDUCE, OP_REDUCE_SCATTER, OP_ALL_GATHER = 0, 1, 2
KERNEL_AUTO, KERNEL_GENERIC, KERNEL_STREAM = 0, 1, 2
KERNEL_NAMES = {"auto": KERNEL_AUTO, "generic": KERNEL_GENERIC, "stream": KERNEL_STREAM}


class AarLaunch(C.Structure):
    """Mirror of ``struct AarLaunch`` — keep field order and types in sync with the header."""

    _fields_ = [
        ("rank", C.c_int32), ("world", C.c_int32), ("emu", C.c_int32),
        ("algo", C.c_int32), ("op", C.c_int32), ("in_dtype", C.c_int32), ("wire_dtype", C.c_int32),
        ("ctas", C.c_int32), ("use_tma", C.c_int32), ("fast_fences", C.c_int32),
        ("role_s", C.c_int32), ("role_c", C.c_int32), ("nvls_unroll", C.c_int32), ("part_kb", C.c_int32), ("part_sc_kb", C.c_int32), ("overlap_flush", C.c_int32), ("oneshot_pull", C.c_int32),
        ("kernel", C.c_int32), ("self_publish", C.c_int32), ("push_chunks", C.c_int32), ("mc_bcast", C.c_int32),
        ("n", C.c_int64), ("max_chunk", C.c_int64), ("os_cap", C.c_int64), ("block_step", C.c_int64),
        ("n_cap", C.c_int64), ("block_step_cap", C.c_int64),
        ("L", C.c_int32), ("slot", C.c_int32),
        ("tag", C.c_uint32),
        ("req_reduce", C.c_int32), ("req_complete", C.c_int32), ("gate_count", C.c_int32),
        ("normalize", C.c_int32),
        ("scale", C.c_float),
        ("skip_dest_mask", C.c_uint32 * MAX_RANKS),
        ("mute_mask", C.c_uint32),
        ("delay_ns", C.c_uint64 * MAX_RANKS),
        ("timeout_ns", C.c_uint64),
        ("src", C.c_void_p * MAX_RANKS),
        ("out", C.c_void_p * MAX_RANKS),
        ("counts", C.c_void_p * MAX_RANKS),
        ("sym", C.c_void_p * MAX_RANKS),
        ("mc_base", C.c_void_p),
        ("scratch", C.c_void_p * MAX_RANKS),
        ("host_flags", C.c_void_p),
        ("trace", C.c_void_p),
    ]


class AarError(RuntimeError):
    pass


_lib = None


def library_path() -> Path:
    return _build.cuda_library_path()


def lib() -> C.CDLL:
    """Loads (building first if needed) the CUDA library and declares its prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.build_cuda()
    l = C.CDLL(str(path))
    i64, i32, vp = C.c_int64, C.c_int, C.c_void_p
    l.aar_sym_bytes.restype = i64
    l.aar_sym_bytes.argtypes = [i64, i64, i32, i32, i32, i64, i64]
    l.aar_sym_flag_bytes.restype = i64
    l.aar_sym_flag_bytes.argtypes = [i64, i64, i32, i32, i32, i64, i64]
    for fn in (l.aar_sym_scat_offset, l.aar_sym_sink_offset):
        fn.restype = i64
        fn.argtypes = [i64, i64, i32, i32, i32, i64, i64]
    l.aar_scratch_bytes.restype = i64
    l.aar_scratch_bytes.argtypes = [i64, i64, i32, i64]
    for fn in (l.aar_scratch_force_offset, l.aar_scratch_dead_offset):
        fn.restype = i64
        fn.argtypes = [i64, i64, i32, i64]
    l.aar_allreduce_round.restype = i32
    l.aar_allreduce_round.argtypes = [C.POINTER(AarLaunch), vp]
    l.aar_allreduce_round_static_scratch.restype = i32
    l.aar_allreduce_round_static_scratch.argtypes = [C.POINTER(AarLaunch), vp]
    l.aar_max_emu_ctas.restype = i32
    l.aar_max_emu_ctas.argtypes = [i32, i32]
    l.aar_max_ctas.restype = i32
    l.aar_max_ctas.argtypes = [i32, i32, i32]
    for fn in (l.aar_round_kernel, l.aar_round_ctas):
        fn.restype = i32
        fn.argtypes = [C.POINTER(AarLaunch)]
    l.aar_heap_create.restype = i32
    l.aar_heap_create.argtypes = [C.POINTER(vp), i32, i32, i32, i64, C.c_char_p, i32]
    l.aar_heap_ptrs.restype = i32
    l.aar_heap_ptrs.argtypes = [vp, C.POINTER(vp), C.POINTER(vp)]
    l.aar_heap_bytes.restype = i64
    l.aar_heap_bytes.argtypes = [vp]
    l.aar_heap_destroy.restype = i32
    l.aar_heap_destroy.argtypes = [vp]
    l.aar_last_error.restype = C.c_char_p
    l.aar_last_error.argtypes = []
    l.aar_abi_sizeof_launch.restype = i32
    l.aar_abi_sizeof_launch.argtypes = []
    if l.aar_abi_sizeof_launch() != C.sizeof(AarLaunch):  # a drifted mirror would corrupt every launch silently
        raise AarError(f"AarLaunch mirror is {C.sizeof(AarLaunch)} bytes, the library's struct {l.aar_abi_sizeof_launch()}: "
                       "ops/cuda.py and include/aar_b200.h are out of sync")
    _lib = l
    return l


def check(rc: int, what: str = "") -> int:
    if rc < 0:
        raise AarError(f"{what or 'libaar_cuda'} failed ({rc}): {lib().aar_last_error().decode()}")
    return rc


def dtype_code(torch_dtype) -> int:
    import torch

    if torch_dtype == torch.float32:
        return F32
    if torch_dtype == torch.bfloat16:
        return BF16
    raise AarError(f"unsupported dtype {torch_dtype}; float32 and bfloat16 are")
