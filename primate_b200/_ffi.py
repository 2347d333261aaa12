"""Thin ctypes binding of the C ABI in include/slq.h (libslq.so, built in-tree by primate_b200._build).

There is NO CPU fallback: if the library is missing it is built with nvcc; if that fails, or if a compute
entry is called without a CUDA device, an exception is raised.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_int, c_int32, c_int64, c_size_t, c_uint64, c_void_p

SLQ_F32, SLQ_F64 = 0, 1
SLQ_OK, SLQ_ERR_ARG, SLQ_ERR_CUDA, SLQ_ERR_UNSUPPORTED, SLQ_ERR_CALLBACK = 0, -1, -2, -3, -4

FUN_IDS = {
	"identity": 0, "abs": 1, "sqrt": 2, "log": 3, "inv": 4, "exp": 5, "smoothstep": 6, "numrank": 7, "softsign": 8,
	None: -1,
}  # fmt: skip

MATMAT_FN = ctypes.CFUNCTYPE(c_int, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p)

_LIB = None


class SlqError(RuntimeError):
	pass


_SIGNATURES = {
	"slq_version": (c_int, []),
	"slq_last_error": (c_char_p, []),
	"slq_device_info": (c_int, [POINTER(c_int), POINTER(c_int), POINTER(c_int)]),
	"slq_launch_count": (ctypes.c_longlong, []),
	"slq_profile_enable": (c_int, [c_int]),
	"slq_profile_read": (c_int, [c_int, POINTER(c_double), POINTER(ctypes.c_longlong), POINTER(ctypes.c_longlong)]),
	"slq_csr_create": (c_int, [c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_int, POINTER(c_void_p)]),
	"slq_csr_create_ex": (c_int, [c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_void_p, POINTER(c_void_p)]),
	"slq_dense_create": (c_int, [c_int64, c_void_p, c_int64, c_int, POINTER(c_void_p)]),
	"slq_dense_create_padded": (c_int, [c_int64, c_int64, c_void_p, c_int64, c_int, POINTER(c_void_p)]),
	"slq_op_set_scratch": (c_int, [c_void_p, c_void_p, c_size_t]),
	"slq_op_set_fused_step": (c_int, [c_void_p, c_int]),
	"slq_dense_tiled_bytes": (c_int, [c_void_p, POINTER(c_size_t)]),
	"slq_dense_set_tiled": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p]),
	"slq_callback_create": (c_int, [c_int64, c_int, MATMAT_FN, c_void_p, POINTER(c_void_p)]),
	"slq_op_destroy": (c_int, [c_void_p]),
	"slq_op_rows": (c_int64, [c_void_p]),
	"slq_op_dtype": (c_int, [c_void_p]),
	"slq_op_nnz": (c_int64, [c_void_p]),
	"slq_op_matmat": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
	"slq_csr_cluster_order": (c_int, [c_int64, c_void_p, c_void_p, c_int, c_void_p]),
	"slq_csr_permute": (c_int, [c_int64, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
	"slq_block_gather_rows": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p]),
	"slq_block_scatter_rows": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p]),
	"slq_probes_rademacher": (c_int, [POINTER(c_uint64), POINTER(c_uint64), c_uint64, c_int, c_int64, c_int, c_void_p, c_void_p]),
	"slq_normals_numpy_version": (c_char_p, []),
	"slq_normals_workspace_bytes": (c_size_t, [c_int64]),
	"slq_normals_scan": (c_int, [POINTER(c_uint64), POINTER(c_uint64), c_uint64, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_size_t,
	                             c_void_p]),
	"slq_probes_normal": (c_int, [POINTER(c_uint64), POINTER(c_uint64), c_uint64, c_int64, c_int, c_int64, c_void_p, c_void_p, c_void_p,
	                              c_int64, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
	"slq_probes_sphere_scale": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p]),
	"slq_pcg64_raw": (c_int, [POINTER(c_uint64), POINTER(c_uint64), c_uint64, c_int64, c_void_p, c_void_p]),
	"slq_block_from_columns": (c_int, [c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p]),
	"slq_block_to_columns": (c_int, [c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p]),
	"slq_lanczos_workspace_bytes": (c_size_t, [c_int64, c_int, c_int, c_int]),
	"slq_lanczos": (c_int, [c_void_p, c_void_p, c_int, c_int, c_double, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
	                        c_void_p, c_void_p, c_size_t, c_void_p]),
	"slq_quadrature": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, POINTER(c_double), c_void_p, c_void_p,
	                           c_void_p, c_void_p, c_void_p, c_void_p]),
	"slq_funm_coeffs": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, POINTER(c_double), c_void_p, c_void_p,
	                            c_void_p, c_void_p, c_void_p, c_void_p]),
	"slq_tqli": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
	"slq_apply_fun": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, POINTER(c_double), c_void_p, c_void_p, c_void_p]),
	"slq_fttr": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
	"slq_stats_update": (c_int, [c_void_p, c_int, c_void_p, c_void_p]),
	"slq_funm_workspace_bytes": (c_size_t, [c_int64, c_int, c_int, c_int]),
	"slq_funm_matmat": (c_int, [c_void_p, c_void_p, c_int, c_int, c_double, c_int, POINTER(c_double), c_void_p, c_void_p, c_size_t, c_void_p]),
	"slq_replay_workspace_bytes": (c_size_t, [c_int64, c_int, c_int, c_int]),
	"slq_lanczos_replay": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
	                               c_size_t, c_void_p]),
	"slq_panel_gram_workspace_bytes": (c_size_t, [c_int, c_int]),
	"slq_panel_gram": (c_int, [c_void_p, c_int, c_int64, c_void_p, c_int, c_int64, c_int64, c_int64, c_int, c_void_p, c_void_p, c_size_t,
	                           c_void_p]),
	"slq_panel_matmul": (c_int, [c_void_p, c_int, c_int64, c_void_p, c_int, c_double, c_double, c_void_p, c_int64, c_int64, c_int, c_void_p]),
	"slq_panel_rowdot": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int, c_int64, c_int, c_void_p, c_void_p]),
	"slq_panel_coldot_workspace_bytes": (c_size_t, [c_int]),
	"slq_panel_coldot": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int, c_int64, c_int64, c_int, c_void_p, c_void_p, c_size_t,
	                             c_void_p]),
	"slq_panel_coldot_static": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int, c_int64, c_int64, c_int, c_void_p, c_void_p]),
	"slq_basis_combine": (c_int, [c_void_p, c_int64, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
}  # fmt: skip

EXPORTED = tuple(_SIGNATURES)


def lib_path() -> str:
	return os.environ.get("SLQ_LIB", os.path.join(os.path.dirname(os.path.abspath(__file__)), "libslq.so"))


def lib() -> ctypes.CDLL:
	"""Loads (building first if needed) libslq.so and declares every entry point of include/slq.h."""
	global _LIB
	if _LIB is None:
		path = lib_path()
		if "SLQ_LIB" not in os.environ:
			from . import _build

			_build.build(verbose=False)  # content-hash check: a stale or missing library is rebuilt, never loaded silently
		L = ctypes.CDLL(path)
		for name, (res, args) in _SIGNATURES.items():
			fn = getattr(L, name)
			fn.restype, fn.argtypes = res, args
		_LIB = L
	return _LIB


def check(rc: int) -> None:
	"""Maps C status codes onto the exception types the reference raises for the same mistakes."""
	if rc == SLQ_OK:
		return
	msg = lib().slq_last_error().decode("utf-8", "replace")
	if rc == SLQ_ERR_ARG:
		raise ValueError(msg)  # pybind11 std::invalid_argument -> ValueError (pylinop.h:24-25)
	if rc == SLQ_ERR_UNSUPPORTED:
		raise NotImplementedError(msg)
	raise SlqError(f"libslq error {rc}: {msg}")


def u128_pair(x: int):
	return (c_uint64 * 2)((x >> 64) & 0xFFFFFFFFFFFFFFFF, x & 0xFFFFFFFFFFFFFFFF)


def fun_params(*vals: float):
	v = list(vals) + [0.0] * (4 - len(vals))
	return (c_double * 4)(*v[:4])
