"""ctypes loader for the two native libraries built by `cme213_sp18_b200.build`.

The device library is the only compute path on a GPU: if it cannot be loaded the ops raise
instead of falling back to PyTorch, so a GPU test can never pass on an eager fallback.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from pathlib import Path

import torch

_LIBDIR = Path(__file__).resolve().parent.parent / "lib"
_lock = threading.Lock()
_dev = None
_host = None

c_void_p, c_int, c_size_t, c_float, c_double = C.c_void_p, C.c_int, C.c_size_t, C.c_float, C.c_double
c_i64, c_u64, c_u32, c_u8 = C.c_int64, C.c_uint64, C.c_uint32, C.c_uint8
P = c_void_p  # every device pointer travels as void*

# name -> argument types (return type is int unless listed in _RESTYPES)
_DEV_SIGS = {
    "cme_version": [],
    "cme_device_info": [P, P, P],
    "cme_l2_fetch_granularity": [c_int, P],
    "cme_shift_bytes": [P, P, c_size_t, c_u8, c_int, P],
    "cme_pagerank_prescale": [P, P, P, c_i64, P],
    "cme_pagerank_step": [P, P, P, P, P, P, P, c_int, c_i64, c_i64, c_i64, P],
    "cme_pagerank_iterate": [P, P, P, P, P, P, P, c_int, c_i64, P],
    "cme_pagerank_gen_edges": [P, c_i64, c_i64, c_i64, c_u64, P],
    "cme_pagerank_block_step": [P, P, P, P, P, P, P, P, c_int, c_i64, c_i64, c_i64, c_int, c_int, P],
    "cme_heat_step": [P, P, c_int, c_int, c_int, c_float, c_float, c_int, c_int, c_float, P],
    "cme_heat_bc": [P, P, c_int, c_int, c_int, c_float, P],
    "cme_heat_slab_step": [P, P, P, P, c_int, c_int, c_int, c_float, c_float, c_float, c_int, c_int, P],
    "cme_sanitize_compact": [P, c_size_t, P, P, P, c_size_t, P],
    "cme_sanitize_compact_noalloc": [P, c_size_t, P, P, P],
    "cme_byte_histogram": [P, c_size_t, c_size_t, c_size_t, P, P],
    "cme_strided_histograms": [P, c_size_t, c_int, P, P],
    "cme_vigenere_shift": [P, P, c_size_t, P, c_int, c_int, P],
    "cme_lag_match_counts": [P, c_size_t, c_int, c_int, P, P],
    "cme_parity_sum": [P, c_size_t, P, P],
    "cme_radix_scratch_bytes": [c_size_t],
    "cme_radix_sort": [P, P, c_size_t, P, c_size_t, P],
    "cme_dgemm": [c_int, c_int, c_int, c_int, c_int, c_double, P, c_int, P, c_int, c_double, P, c_int, P],
    "cme_dgemm_simt": [c_int, c_int, c_int, c_int, c_int, c_double, P, c_int, P, c_int, c_double, P, c_int, P],
    "cme_d_bias_sigmoid": [P, P, c_int, c_int, P],
    "cme_d_bias_softmax_ce": [P, P, P, P, c_double, P, c_int, c_int, P],
    "cme_d_dsigmoid": [P, P, P, c_size_t, P],
    "cme_d_row_sum": [P, P, c_int, c_int, P],
    "cme_d_sgd": [P, P, c_size_t, c_double, c_double, P],
    "cme_d_axpby": [P, P, P, c_double, c_double, c_size_t, P],
    "cme_u8_to_bf16_rows": [P, P, c_i64, c_int, c_int, c_int, P],
    "cme_mlp_fwd1": [P, P, P, P, c_i64, c_int, c_int, c_int, c_int, P],
    "cme_mlp_fwd2_softmax": [P, P, P, P, P, P, P, P, c_float, c_i64, c_int, c_int, c_int, P],
    "cme_mlp_bwd_elem": [P, P, P, P, P, c_i64, c_int, c_int, c_int, P],
    "cme_mlp_fwd2_tc": [P, P, P, P, P, P, P, P, P, c_float, c_i64, c_int, c_int, c_int, P],
    "cme_mlp_bwd_dz1_tc": [P, P, P, P, c_i64, c_int, c_int, P],
    "cme_mlp_bwd_tail_tc": [P, P, P, P, P, c_int, c_i64, c_int, c_int, c_int, P],
    "cme_mlp_recast_w2": [P, P, c_int, P, c_int, c_int, P],
    "cme_mlp_dw2_tc": [P, P, P, c_i64, c_int, c_int, c_int, P],
    "cme_f32_to_bf16_2d": [P, c_int, P, c_int, c_int, c_int, P],
    "cme_mlp_dw1": [P, P, P, c_i64, c_int, c_int, c_int, c_int, c_int, P],
    "cme_gemm_bf16": [P, P, P, c_i64, c_int, c_int, c_int, c_int, c_int, P],
    "cme_allreduce_sgd": [P, P, P, c_int, c_int, c_i64, c_i64, c_int, c_int, c_i64, c_i64, c_float, c_float, c_i64, c_i64, P],
    "cme_allreduce_sgd_mc": [P, P, P, P, c_int, c_int, c_i64, c_i64, c_int, c_int, c_i64, c_i64, c_float, c_float, c_i64, c_i64, P],
    "cme_allreduce_sgd_f64": [P, P, P, c_int, c_int, c_i64, c_i64, c_i64, c_i64, c_double, c_double, c_int, P],
    "cme_mlp_stage_recast": [P, P, P, c_i64, c_int, c_int, P, P, c_int, P, c_int, c_int, P],
    "cme_update_slice": [c_i64, c_i64, c_int, c_int, P],
    "cme_mlp_dw1_scratch_bytes": [],
    "cme_mlp_dw1_plan_query": [c_i64, c_int, c_int, P],
    "cme_mlp_bwd_tail_scratch_floats": [c_int],
    "cme_mlp_bwd_tail_det": [P, P, P, P, P, c_int, c_i64, c_int, c_int, c_int, P, P],
    "cme_mlp_fwd_fused_ok": [c_i64, c_int],
    "cme_mlp_fwd_scratch_floats": [c_i64],
    "cme_mlp_fwd_fused": [P, P, P, P, P, P, P, P, P, P, P, P, P, c_i64, c_float, c_i64, c_int, c_int, c_int, c_int, c_int, P],
    "cme_mlp_dw1_sync": [P, P, P, c_i64, c_int, c_int, c_int, c_int, c_int, P, P, c_int, c_int, P, P],
    "cme_mlp_update_stream": [P, c_float, c_float, P],
    "cme_mlp_dw1_update": [P, c_i64, c_float, c_float, c_int, P],
    "cme_mlp_status": [P, P],
    "cme_mlp_unit_slice": [P, c_int, c_int, P],
    "cme_mlp_set_step_args": [P, P, P, c_float, c_float, c_float, P],
    "cme_mlp_step_ex": [P, P, P, c_i64, c_float, c_float, c_float, P, P, P, P],
    "cme_mlp_step": [P, P, P, c_i64, c_float, c_float, c_float, P, P],
    "cme_mlp_step_launches": [P],
    "cme_mlp_layout": [P, P],
    "cme_peer_barrier": [P, c_int, c_int, c_int, c_u32, c_int, c_int, P],
    "cme_peer_copy": [P, P, c_size_t, P],
    "cme_peer_copy_bg": [P, P, c_size_t, c_int, P],
    "cme_peer_memcpy": [P, P, c_size_t, P],
}
_DEV_RESTYPES = {"cme_radix_scratch_bytes": c_size_t, "cme_mlp_dw1_scratch_bytes": c_size_t,
                 "cme_mlp_fwd_scratch_floats": c_size_t, "cme_mlp_bwd_tail_scratch_floats": c_size_t}

_HOST_SIGS = {
    "cmeh_num_threads": [],
    "cmeh_mlp_train": [P, P, c_i64, c_int, c_int, c_int, P, P, P, P, c_double, c_double, c_int, c_int, c_int, P, c_i64],
    "cmeh_mlp_forward": [P, c_i64, c_int, c_int, c_int, P, P, P, P, P, P],
    "cmeh_parity_sum_serial": [P, c_size_t, P, P],
    "cmeh_parity_sum_omp": [P, c_size_t, c_int, P, P],
    "cmeh_radix_block_histograms": [P, c_size_t, c_u32, c_u32, c_u32, c_size_t, P],
    "cmeh_radix_reduce_histograms": [P, c_u32, c_u32, P],
    "cmeh_radix_scan_global": [P, c_u32, P],
    "cmeh_radix_block_exscan": [P, P, c_u32, c_u32, P],
    "cmeh_radix_populate": [P, c_u32, c_u32, c_u32, c_size_t, P, c_size_t, P],
    "cmeh_radix_sort_serial": [P, P, c_size_t, c_u32],
    "cmeh_radix_sort_parallel": [P, P, c_size_t, c_u32, c_u32, c_int],
    "cmeh_std_sort": [P, c_size_t],
    "cmeh_shift_bytes": [P, P, c_size_t, c_u8],
    "cmeh_pagerank": [P, P, P, P, P, c_int, c_i64],
    "cmeh_pagerank_graph": [c_i64, c_int, c_u64, P, P, P],
    "cmeh_heat_cpu": [P, P, c_int, c_int, c_int, c_int, c_float, c_float, c_float, c_int, c_int],
    "cmeh_heat_init": [P, c_int, c_int, c_double, c_double],
    "cmeh_max_ulp_f32": [P, P, c_size_t, P],
    "cmeh_max_ulp_f64": [P, P, c_size_t, P],
    "cmeh_letter_histogram": [P, c_size_t, c_size_t, c_size_t, P],
    "cmeh_idx_info": [C.c_char_p, P, P],
    "cmeh_idx_read": [C.c_char_p, c_i64, c_i64, P],
}
_HOST_RESTYPES = {"cmeh_mlp_train": c_i64, "cmeh_max_ulp_f32": c_i64, "cmeh_max_ulp_f64": c_i64, "cmeh_idx_info": c_i64,
                  "cmeh_idx_read": c_i64}


def _declare(lib, sigs, restypes, strict):
    for name, args in sigs.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            if strict:
                raise
            continue
        fn.argtypes = args
        fn.restype = restypes.get(name, c_int)


def _load(path: Path):
    if not path.exists():
        # build on demand (a fresh checkout, or the driver's build() not run yet)
        from .. import build as _build

        _build.build()
    return C.CDLL(str(path))


def host_lib():
    global _host
    with _lock:
        if _host is None:
            _host = _load(_LIBDIR / "libcme213b_host.so")
            _declare(_host, _HOST_SIGS, _HOST_RESTYPES, strict=True)
    return _host


def device_lib():
    global _dev
    with _lock:
        if _dev is None:
            lib = _load(_LIBDIR / "libcme213b.so")
            _declare(lib, _DEV_SIGS, _DEV_RESTYPES, strict=False)
            lib.cme_last_error.restype = C.c_char_p
            lib.cme_last_error.argtypes = []
            _dev = lib
    return _dev


def device_lib_path() -> str:
    return str(_LIBDIR / "libcme213b.so")


class NativeError(RuntimeError):
    pass


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = device_lib().cme_last_error().decode(errors="replace")
        raise NativeError(f"{what or 'native call'} failed (rc={rc}): {msg}")


def stream_ptr(stream: torch.cuda.Stream | None = None) -> int:
    s = stream if stream is not None else torch.cuda.current_stream()
    return int(s.cuda_stream)


def ptr(t: torch.Tensor | None) -> int | None:
    """Device (or host) address of a tensor; None passes NULL."""
    if t is None:
        return None
    if isinstance(t, int):          # a raw (possibly offset or peer-mapped) address
        return t
    return int(t.data_ptr())


def require_cuda(*tensors: torch.Tensor) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise ValueError("this op runs on the GPU only; got a CPU tensor")


def np_ptr(a) -> int:
    return int(a.ctypes.data)


def env_flag(name: str, default: bool = False) -> bool:
    v = os.environ.get(name)
    if v is None:
        return default
    return v.strip().lower() not in ("", "0", "false", "no", "off")
