"""ctypes binding of libtmog_b200.so -- the same C ABI a JNI shim binds (include/tmog_b200.h).

There is no fallback: if the library has not been built (``python -c "import __graft_entry__ as g; g.build()"``) the
import of anything that computes fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os
import re
from typing import Dict, List

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TMOG_B200_LIB") or os.path.join(_HERE, "libtmog_b200.so")   # override: A/B kernel builds
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "tmog_b200.h")

REAL_NSTAT = 8


class TmogError(RuntimeError):
    """A call into libtmog_b200 failed; ``code`` is the TMOG_E* status."""

    def __init__(self, code: int, message: str):
        super().__init__(f"libtmog_b200: {message} (status {code})")
        self.code = code
        self.message = message


class TmogInvalidArgument(TmogError, ValueError):
    """TMOG_EINVAL -- the reference would have thrown IllegalArgumentException (a failed ``require``)."""


_lib = None

_p = C.c_void_p
_i64 = C.c_int64
_int = C.c_int
_dbl = C.c_double
_sz = C.c_size_t
_u64 = C.c_uint64

# name -> argtypes; every function returns int except the two noted below
_SIGNATURES: Dict[str, List] = {
    "tmog_real_pilot_shift_dev": [_p, _i64, _p, _i64, _p, _i64, _int, _p, _p],
    "tmog_real_fit_scratch_bytes": [_int, C.POINTER(_sz)],
    "tmog_real_fit_partials_dev": [_p, _i64, _p, _i64, _p, _i64, _int, _p, _p, _sz, _p, _p],
    "tmog_real_fit_partials_noalloc_dev": [_p, _i64, _p, _i64, _p, _i64, _int, _p, _p, _p],
    "tmog_real_partials_merge": [_p, _p, _int],
    "tmog_real_finish": [_p, _p, _int, _int, _dbl, _int, _p, _p, _p, _p, _p, _p, _p, _p],
    "tmog_real_transform_dev": [_p, _i64, _p, _i64, _i64, _int, _p, _int, _p, _i64, _i64, _p],
    "tmog_binary_transform_dev": [_p, _i64, _p, _i64, _i64, _int, _int, _int, _p, _i64, _i64, _p],
    "tmog_clean_strings": [_p, _p, _i64, _p, _i64, _p],
    "tmog_dedupe_strings": [_p, _p, _i64, _p, _p],
    "tmog_pivot_topk": [_p, _p, _p, _i64, _int, _i64, _p, _p],
    "tmog_pivot_column_finish": [_p, _p, _i64, _int, _p, _int, _int, _i64, _p, _p, _p, _p, _i64, _p],
    "tmog_hash_bytes": [_p, _i64, C.c_uint64, _p],
    "tmog_pivot_dict_prepare": [_p, _p, _i64, _int, _p],
    "tmog_pivot_dict_finish": [_p, _p, _int, _int, _i64, _p, _p, _p, _p, _i64, _p],
    "tmog_pivot_dict_free": [_p],
    "tmog_pivot_column_map": [_p, _p, _i64, _int, _p, _p, _int, _p],
    "tmog_pivot_count_dev": [_p, _i64, _i64, _int, _p, _int, _p, _p, _i64, _p, _p, _p],
    "tmog_binary_codes_dev": [_p, _i64, _p, _i64, _i64, _int, _p, _i64, _p],
    "tmog_pivot_transform_dev": [_p, _i64, _i64, _int, _p, _p, _p, _i64, _i64, _p],
    "tmog_set_count_dev": [_p, _p, _i64, _p, _int, _int, _p, _int, _p, _p, _p],
    "tmog_set_stats_dev": [_p, _p, _i64, _p, _int, _int, _int, _p, _int, _p, _p, _p, _p, _p, _dbl, _p, _p, _p],
    "tmog_pivot_label_sum_dev": [_p, _i64, _i64, _int, _p, _p, _p, _p, _dbl, _p, _p],
    "tmog_set_transform_dev": [_p, _p, _i64, _p, _int, _int, _int, _p, _p, _p, _p, _p, _p, _p],
    "tmog_text_stats_layout": [_int, _int, _int, C.POINTER(_sz), C.POINTER(_sz)],
    "tmog_text_stats_table_bytes": [_int, _int, C.POINTER(_sz)],
    "tmog_text_stats_dev": [_p, _int, _i64, _p, _int, _int, _int, _int, _p, _p, _p],
    "tmog_text_stats_values_dev": [_p, _int, _i64, _p, _int, _int, _int, _int, _p, _p, _p],
    "tmog_text_length_hist_dev": [_p, _int, _i64, _i64, _int, _int, _int, _int, _p, _p, _p, _p],
    "tmog_text_codes_dev": [_p, _p, _int, _i64, _int, _int, _p, _p, _i64, _p, _p],
    "tmog_text_hash_scratch_bytes": [_i64, C.POINTER(_sz)],
    "tmog_text_hash_scratch_bytes2": [_i64, _int, _p, _int, C.POINTER(_sz)],
    "tmog_text_hash_stats_dev": [_p, _int, _i64, _p, _p, _p, _p, _p, _p, _p, _p, _p, _dbl, _p, _p, _p, _sz, _p],
    "tmog_text_hash_transform_dev": [_p, _int, _i64, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i64, _p, _p, _sz, _p],
    "tmog_text_hash_fit_transform_dev": [_p, _int, _i64, _p, _p, _p, _p, _p, _p, _p, _p, _p, _dbl, _p, _p, _p, _p, _p, _p,
                                         _p, _i64, _p, _p, _sz, _p],
    "tmog_text_cells_pack_bytes": [_i64, _int, _int, C.POINTER(_sz)],
    "tmog_text_hash_stats_pack_dev": [_p, _int, _i64, _p, _p, _p, _p, _p, _p, _p, _p, _p, _dbl, _p, _p, _p, _sz, _p, _p,
                                      _u64, _p, _p, _p],
    "tmog_text_hash_transform_packed_dev": [_int, _i64, _p, _p, _p, _u64, _p, _p, _p, _p, _p, _p, _i64, _p, _p],
    "tmog_dense_split_dev": [_p, _i64, _i64, _int, _p, _p, _i64, _p, _i64, _p, _p],
    "tmog_indices_narrow_dev": [_p, _i64, _p, _p, _p, _p],
    "tmog_values_narrow_dev": [_p, _i64, _p, _p, _p, _p, _int, _p, _p],
    "tmog_sparse_rows_select_dev": [_p, _p, _p, _p, _i64, _p, _p, _p, _p, _p, _p],
    "tmog_label_classes_scratch_bytes": [C.POINTER(_sz)],
    "tmog_label_classes_dev": [_p, _i64, _int, _p, _p, _p, _p, _p, _sz, _p],
    "tmog_label_class_ids_dev": [_p, _i64, _p, _p, _int, _p, _p, _p],
    "tmog_null_class_counts_dev": [_p, _i64, _p, _i64, _int, _int, _int, _p, _p],
    "tmog_dense_pilot_shift_dev": [_p, _i64, _i64, _int, _p, _p, _p],
    "tmog_dense_stats_scratch_bytes": [_i64, _int, C.POINTER(_sz)],
    "tmog_dense_stats_dev": [_p, _i64, _i64, _int, _p, _p, _int, _p, _p, _p, _sz, _p, _p, _p],
    "tmog_dense_gather_dev": [_p, _i64, _i64, _p, _int, _p, _i64, _p],
    "tmog_dense_gram_dev": [_p, _i64, _i64, _int, _p, _p, _p, _p],
    "tmog_onehot_densify_dev": [_p, _i64, _i64, _int, _p, _i64, _i64, _p],
    "tmog_sparse_rows_densify_dev": [_p, _p, _p, _p, _i64, _p, _i64, _i64, _p],
    "tmog_pearson_from_gram": [_int, _dbl, _p, _p, _p],
    "tmog_contingency_stats": [_p, _int, _int, _p, _p, _p, _p, _p, _p],
    "tmog_contingency_stats_batch": [_int, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p],
    "tmog_pearson_from_cov": [_int, _dbl, _p, _p, _dbl, _p],
    "tmog_bucket_stats": [_i64, _int, _p, _p, _p, _p, _p, _dbl, _dbl, _p, _p, _p, _p, _p],
    "tmog_sanity_decide": [_int, _p, _p, _p, _int, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p],
    "tmog_real_transform_select_dev": [_p, _i64, _p, _i64, _i64, _int, _p, _int, _p, _int, _p, _i64, _i64, _p],
    "tmog_real_finish_with_fills": [_p, _p, _int, _p, _int, _p, _p, _p, _p, _p, _p, _p],
    "tmog_integral_mode_scratch_bytes": [_int, _i64, _p],
    "tmog_integral_mode_layout": [_int, _i64, _p],
    "tmog_integral_mode_dev": [_p, _i64, _p, _i64, _i64, _int, _i64, _p, _sz, _p, _p, _p],
    "tmog_integral_pilot_shift_dev": [_p, _i64, _p, _i64, _p, _i64, _int, _p, _p],
    "tmog_integral_fit_partials_dev": [_p, _i64, _p, _i64, _p, _i64, _int, _p, _p, _sz, _p, _p],
    "tmog_integral_transform_dev": [_p, _i64, _p, _i64, _i64, _int, _p, _int, _p, _i64, _i64, _p],
    "tmog_integral_transform_select_dev": [_p, _i64, _p, _i64, _i64, _int, _p, _int, _p, _int, _p, _i64, _i64, _p],
    "tmog_binary_transform_select_dev": [_p, _i64, _p, _i64, _i64, _int, _int, _int, _p, _int, _p, _i64, _i64, _p],
    "tmog_ctx_create": [_int, C.POINTER(_p)],
    "tmog_ctx_destroy": [_p],
    "tmog_pinned_alloc": [_sz, C.POINTER(_p)],
    "tmog_pinned_free": [_p],
    "tmog_pivot_fit_host": [_p, _p, _i64, _i64, _int, _p, _p, _p, _p, _i64, _int, _int, _i64, _int, _p, _p, _p, _p, _p, _i64,
                            _p, _p, _p],
    "tmog_pivot_transform_host": [_p, _p, _i64, _i64, _int, _p, _p, _p, _int, _p, _i64, _p, _i64, _i64],
    "tmog_text_hash_stats_host": [_p, _int, _i64, _p, _p, _p, _p, _p, _p, _p, _int, _p, _p, _p, _p, _p, _p, _p, _p],
    "tmog_text_hash_transform_host": [_p, _int, _i64, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i64, _p],
    "tmog_vector_export_count_dev": [_p, _int, _i64, _int, _p, _p, _p, _p, _p],
    "tmog_vector_export_write_dev": [_p, _int, _i64, _int, _p, _p, _p, _p, _p, _p],
    "tmog_rff_numeric_summary_dev": [_p, _int, _i64, _p, _i64, _i64, _int, _p, _p],
    "tmog_rff_numeric_hist_dev": [_p, _int, _i64, _p, _i64, _i64, _int, _int, _p, _p, _p, _p, _p, _p, _p],
    "tmog_rff_rows_hist_dev": [_p, _p, _p, _p, _i64, _int, _int, _p, _p, _p, _p, _p],
    "tmog_rff_code_count_dev": [_p, _i64, _i64, _int, _p, _p, _p, _p, _p, _p, _dbl, _p, _p],
    "tmog_rff_null_label_sums_dev": [_p, _int, _i64, _p, _p, _dbl, _p, _p],
    "tmog_rff_numeric_splits": [_dbl, _dbl, _int, _p],
    "tmog_rff_js_divergence": [_p, _p, _int, _p],
    "tmog_rff_null_label_corr": [_dbl, _dbl, _dbl, _int, _p, _p, _p],
    "tmog_comm_unique_id": [_p],
    "tmog_comm_init": [_p, _int, _int, _int, C.POINTER(_p)],
    "tmog_comm_destroy": [_p],
    "tmog_comm_allreduce_dev": [_p, _p, _i64, _int, _int, _p],
    "tmog_comm_allreduce_host": [_p, _p, _i64, _int, _int],
    "tmog_real_fit_host": [_p, _p, _i64, _p, _i64, _p, _i64, _int, _i64, _int, _p, _p, _p, _int, _p, _p, _p, _p],
    "tmog_real_transform_host": [_p, _p, _i64, _p, _i64, _i64, _int, _p, _int, _i64, _p, _int, _p, _i64, _i64],
}


def declared_symbols() -> List[str]:
    """Every function the public header declares (used by the export test)."""
    with open(HEADER_PATH) as f:
        src = f.read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(tmog_[a-z0-9_]+)\s*\(", src)))


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.tmog_last_error.restype = C.c_char_p
        L.tmog_last_error.argtypes = []
        L.tmog_abi_version.restype = _int
        L.tmog_abi_version.argtypes = []
        for name, args in _SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = _int
            fn.argtypes = args
        _lib = L
    return _lib


def check(status: int) -> None:
    if status != 0:
        msg = lib().tmog_last_error().decode("utf-8", "replace")
        if status == -1:
            raise TmogInvalidArgument(status, msg)
        raise TmogError(status, msg)


DEV_CALLS = [0]      # calls of *_dev entry points so far (each one queues at least one of the library's kernels)


def call(name: str, *args) -> None:
    if name.endswith("_dev"):
        DEV_CALLS[0] += 1
    check(getattr(lib(), name)(*args))
