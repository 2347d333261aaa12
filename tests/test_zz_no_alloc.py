"""slq_panel_coldot_static: the allocation-free variant of slq_panel_coldot (per-CTA partial table in a static __device__
array of the library instead of a caller workspace; include/slq.h).  No counterpart in the reference -- added for callers
that cannot allocate device memory.  Called through the C ABI and checked against the existing entry point on the same
panels, on two calls in a row that share the static array (the second with a different shape, so stale contents of the
first would show), plus the documented size limit.  Collected last (file name) so that a failure here hides nothing."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _coldot_both(X, Y, row0, row1, code):
	"""(static-scratch result, caller-workspace result) of the column dots of two device panels, both through ctypes."""
	import torch

	from primate_b200 import _ffi, engine

	L = _ffi.lib()
	m = int(X.shape[1])
	out_s = torch.full((m,), float("nan"), dtype=torch.float64, device="cuda")
	out_w = torch.full((m,), float("nan"), dtype=torch.float64, device="cuda")
	ws = torch.empty(max(int(L.slq_panel_coldot_workspace_bytes(m)), 1), dtype=torch.uint8, device="cuda")
	st = engine._stream_ptr()
	_ffi.check(L.slq_panel_coldot_static(X.data_ptr(), int(X.stride(0)), Y.data_ptr(), int(Y.stride(0)), m, row0, row1, code,
	                                     out_s.data_ptr(), st))  # fmt: skip
	_ffi.check(L.slq_panel_coldot(X.data_ptr(), int(X.stride(0)), Y.data_ptr(), int(Y.stride(0)), m, row0, row1, code, out_w.data_ptr(),
	                              ws.data_ptr(), ws.numel(), st))  # fmt: skip
	torch.cuda.synchronize()
	return out_s.cpu().numpy(), out_w.cpu().numpy()


def test_static_scratch_coldot_matches_workspace_coldot_on_two_calls_in_a_row():
	import torch

	from primate_b200 import _ffi

	assert torch.cuda.is_available(), "GPU tests need a CUDA device"
	rng = np.random.default_rng(0)
	# first call: wide panel, many CTAs (fills most of the static table); second: narrower, fewer rows, a row range, fp32
	cases = [(40000, 256, 0, 40000, np.float64), (3001, 37, 100, 2900, np.float32), (40000, 96, 0, 40000, np.float64), (5, 1, 2, 2, np.float64)]
	for n, m, row0, row1, dtype in cases:
		Xh, Yh = rng.standard_normal((n, m)).astype(dtype), rng.standard_normal((n, m)).astype(dtype)
		X, Y = torch.from_numpy(Xh).cuda(), torch.from_numpy(Yh).cuda()
		code = _ffi.SLQ_F64 if dtype == np.float64 else _ffi.SLQ_F32
		got, ref = _coldot_both(X, Y, row0, row1, code)
		exact = np.einsum("ij,ij->j", Xh[row0:row1].astype(np.float64), Yh[row0:row1].astype(np.float64))
		tol = (1e-12 if dtype == np.float64 else 1e-5) * max(n, 1)  # the tolerance of tests/test_gpu_panels.py
		assert np.all(np.isfinite(got)) and np.max(np.abs(got - ref)) <= tol, (n, m)
		assert np.array_equal(got, ref), (n, m)  # same kernels, same grid, same summation order
		assert np.max(np.abs(got - exact)) <= tol, (n, m)
		again, _ = _coldot_both(X, Y, row0, row1, code)
		assert np.array_equal(again, got)  # deterministic, whatever the previous call left in the array


def test_static_scratch_coldot_rejects_panels_wider_than_its_limit():
	import torch

	from primate_b200 import _ffi

	L = _ffi.lib()
	m = 257  # SLQ_COLDOT_STATIC_MAX_M + 1
	X = torch.ones((64, m), dtype=torch.float64, device="cuda")
	out = torch.zeros(m, dtype=torch.float64, device="cuda")
	rc = L.slq_panel_coldot_static(X.data_ptr(), m, X.data_ptr(), m, m, 0, 64, _ffi.SLQ_F64, out.data_ptr(), None)
	assert rc == -1  # SLQ_ERR_ARG
	msg = L.slq_last_error().decode()
	assert "256" in msg and "slq_panel_coldot" in msg
	assert float(out.sum().item()) == 0.0  # nothing was launched
