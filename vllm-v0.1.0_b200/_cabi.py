"""ctypes binding of libpaged_b200.so (the C ABI in include/paged_attention_b200.h).

There is no fallback: if the library is missing, `lib()` raises, and every
operator in attention_ops / cache_ops fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libpaged_b200.so")

PA_ABI_VERSION = 2          # include/paged_attention_b200.h; lib() refuses a library that reports another
PA_F16, PA_BF16, PA_F32 = 0, 1, 2
PA_COPY_D2D, PA_COPY_D2H, PA_COPY_H2D = 0, 1, 2
PA_SLOTS_AUTO, PA_SLOTS_SCATTERED, PA_SLOTS_RUNS = 0, 1, 2

PA_OK = 0
PA_ERR_UNSUPPORTED_HEAD_SIZE = -1
PA_ERR_UNSUPPORTED_BLOCK_SIZE = -2
PA_ERR_UNSUPPORTED_DTYPE = -3
PA_ERR_INVALID_ARGUMENT = -4
PA_ERR_CUDA = -5
PA_ERR_NEED_TABLE = -6


class DecodeTuning(C.Structure):
    """Mirror of `pa_decode_tuning`."""
    _fields_ = [("variant", C.c_int32), ("num_warps", C.c_int32),
                ("num_stages", C.c_int32), ("num_splits", C.c_int32)]


_vp, _i32, _i64, _f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float


class StepBatch(C.Structure):
    """Mirror of `pa_step_batch` (host arrays of one step)."""
    _fields_ = [("num_seqs", _i32), ("seq_ids", _vp), ("last_tokens", _vp), ("context_lens", _vp),
                ("num_blocks", _vp), ("last_block", _vp), ("table_offsets", _vp), ("table_blocks", _vp),
                ("num_prompts", _i32), ("prompt_offsets", _vp), ("prompt_tokens", _vp),
                ("prompt_table_offsets", _vp), ("prompt_table_blocks", _vp)]


class StepOutputs(C.Structure):
    """Mirror of `pa_step_outputs` (device pointers)."""
    _fields_ = [("tokens", _vp), ("positions", _vp), ("slot_mapping", _vp), ("context_lens", _vp),
                ("block_tables", _vp), ("block_table_width", _i32)]


class StepInfo(C.Structure):
    """Mirror of `pa_step_info`."""
    _fields_ = [("max_context_len", _i32), ("max_num_blocks", _i32), ("num_tokens", _i32),
                ("num_incremental", _i32), ("num_full", _i32), ("num_diffed", _i32), ("h2d_bytes", _i64)]


# name -> (restype, argtypes); must list every PA_API symbol of the header
SIGNATURES = {
    "pa_abi_version": (C.c_int, []),
    "pa_last_error": (C.c_char_p, []),
    "pa_compiled_sm": (C.c_int, []),
    "pa_launch_count": (C.c_uint64, []),
    "pa_release_workspaces": (C.c_int, []),
    "pa_stream_read_probe": (C.c_int, [_vp, _i64, _i32, _vp, _vp]),
    "pa_decode_attention": (C.c_int, [_vp, _vp, _vp, _vp, _f32, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32,
                                      _i64, _i32, _vp]),
    "pa_decode_attention_ex": (C.c_int, [_vp, _vp, _vp, _vp, _f32, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32,
                                         _i64, _i32, C.POINTER(DecodeTuning), _vp]),
    "pa_decode_attention_static": (C.c_int, [_vp, _vp, _vp, _vp, _f32, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32,
                                             _i64, _i32, C.POINTER(DecodeTuning), _vp]),
    "pa_decode_attention_reserve": (C.c_int, [_i32, _i32, _i32, _i32, _i32, _i32, _i32, C.POINTER(DecodeTuning), _vp]),
    "pa_decode_attention_plan": (C.c_int, [_i32, _i32, _i32, _i32, _i32, _i32, C.POINTER(DecodeTuning)]),
    "pa_prefill_attention": (C.c_int, [_vp, _vp, _vp, _vp, C.POINTER(_i32), _i32, _i32, _i32, _i64, _i64, _i64, _i64,
                                       _f32, _i32, _vp]),
    "pa_prefill_attention_debug": (C.c_int, [_vp, _vp, _vp, _vp, C.POINTER(_i32), _i32, _i32, _i32, _i64, _i64, _i64,
                                             _i64, _f32, _i32, _vp, _vp]),
    "pa_multi_query_cached_kv_workspace_bytes": (_i64, [_i64, _i32, _i32, _i32]),
    "pa_multi_query_cached_kv_attention": (C.c_int, [_vp, _vp, _vp, _vp, _f32, _vp, C.POINTER(_i32), C.POINTER(_i32),
                                                     _i32, _i32, _i32, _i32, _i32, _i64, _i64, _i32, _vp, _i64, _vp]),
    "pa_reshape_and_cache": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i64, _i64, _i32, _vp]),
    "pa_reshape_and_cache_ex": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i64, _i64, _i32, _i32, _vp]),
    "pa_gather_cached_kv": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i64, _i64, _i32, _vp]),
    "pa_rotary_embedding_neox": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i64, _i64, _i32, _vp]),
    "pa_rope_reshape_and_cache": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32,
                                            _i64, _i64, _i64, _i32, _vp]),
    "pa_allreduce": (C.c_int, [C.POINTER(_vp), C.POINTER(_vp), _vp, _i64, _i64, _i32, _i32, _i32, _vp]),
    "pa_allreduce_flag_bytes": (C.c_int, []),
    "pa_allreduce_emulated": (C.c_int, [C.POINTER(_vp), C.POINTER(_vp), _i64, _i64, _i32, _i32, _vp]),
    "pa_step_state_create": (C.c_int, [_i32, _i32, _i32, C.POINTER(_vp)]),
    "pa_step_state_destroy": (C.c_int, [_vp]),
    "pa_step_state_reset": (C.c_int, [_vp]),
    "pa_step_prepare_inputs": (C.c_int, [_vp, C.POINTER(StepBatch), C.POINTER(StepOutputs), C.POINTER(StepInfo),
                                         _vp, _vp]),
    "pa_copy_blocks": (C.c_int, [C.POINTER(_vp), C.POINTER(_vp), _i32, C.POINTER(_i64), _i64, _i64, _vp]),
    "pa_swap_blocks": (C.c_int, [_vp, _vp, _i64, C.POINTER(_i64), C.POINTER(_i64), _i64, _i32, _vp]),
    "pa_swap_blocks_ex": (C.c_int, [_vp, _vp, _i64, C.POINTER(_i64), C.POINTER(_i64), _i64, _i32, _i32, _vp]),
}

_lib = None
_lock = threading.Lock()


def lib() -> C.CDLL:
    """Loads the shared library once; raises RuntimeError if it has not been built."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise RuntimeError(
                        f"{LIB_PATH} is missing: build it with `python vllm-v0.1.0_b200/build.py` "
                        "(there is no CPU or PyTorch fallback for these operators)")
                handle = C.CDLL(LIB_PATH)
                handle.pa_abi_version.restype = C.c_int
                found = handle.pa_abi_version()
                if found != PA_ABI_VERSION:
                    raise RuntimeError(f"{LIB_PATH} reports C ABI version {found}, this binding needs {PA_ABI_VERSION}: "
                                       "rebuild it with `python vllm-v0.1.0_b200/build.py`")
                for name, (res, args) in SIGNATURES.items():
                    fn = getattr(handle, name)
                    fn.restype = res
                    fn.argtypes = args
                _lib = handle
    return _lib


TORCH_OPS_PATH = os.path.join(_HERE, "lib", "pa_torch_ops.so")
_torch_ops = None
_binding = os.environ.get("PA_BINDING", "")      # "" = compiled binding when built, else ctypes


def torch_ops():
    """The compiled operator binding (csrc/torch_binding.cpp -> lib/pa_torch_ops.so), or None when it is
    not used.  Both bindings end in the same C ABI; this one unpacks tensors in C++ (~2 us per call)."""
    global _torch_ops
    if _binding == "ctypes":
        return None
    if _torch_ops is None:
        if not os.path.exists(TORCH_OPS_PATH):
            if _binding == "torch":
                raise RuntimeError(f"{TORCH_OPS_PATH} is missing: build it with `python vllm-v0.1.0_b200/build.py`")
            return None
        import importlib.util
        lib()                                       # libpaged_b200.so first: fails loudly if it is missing
        spec = importlib.util.spec_from_file_location("pa_torch_ops", TORCH_OPS_PATH)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _torch_ops = mod
    return _torch_ops


def use_binding(name: str) -> None:
    """"torch", "ctypes" or "" (automatic).  Tests run both."""
    global _binding
    if name not in ("", "torch", "ctypes"):
        raise ValueError(name)
    _binding = name


def last_error() -> str:
    return lib().pa_last_error().decode("utf-8", "replace")


def check(rc: int) -> None:
    """Maps a non-zero status to RuntimeError, as TORCH_CHECK does in the reference."""
    if rc != PA_OK:
        raise RuntimeError(f"libpaged_b200: {last_error()} (status {rc})")


def launch_count() -> int:
    return int(lib().pa_launch_count())
