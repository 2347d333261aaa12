"""The band-major copy of a dense fp64 operator (libslq: slq_dense_tiled_bytes / slq_dense_set_tiled) feeds the tensor-core
product kernel the same stage images as the row-major operator does, so Y = A X must be BIT-identical with and without
it -- for every block width, CTA height (SLQ_DENSE_RPT picks are by n), padded orders and the split-K path of small n."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n", [16, 150, 1000, 1008, 2464, 4096, 5000])
def test_tiled_copy_is_bit_identical_and_matches_numpy(n):
	import torch

	from primate_b200 import engine

	rng = np.random.default_rng(n)
	B = rng.standard_normal((n, n))
	A = (B + B.T) / 2
	plain = engine.DeviceOperator(A, tiled=False)
	tiled = engine.DeviceOperator(A, tiled=True)
	assert tiled.tiled and not plain.tiled
	for k in (8, 16, 32, 64, 1, 4):
		X = rng.standard_normal((n, k))
		Xd = torch.from_numpy(X).cuda()
		Y0, Y1 = plain.matmat(Xd), tiled.matmat(Xd)
		assert torch.equal(Y0, Y1), (n, k)
		assert np.max(np.abs(Y1.cpu().numpy() - A @ X)) < 1e-12 * n


def test_tiled_copy_declined_for_fp32_and_reported_size():
	import ctypes

	from primate_b200 import _ffi, engine

	A = np.eye(64, dtype=np.float32)
	op = engine.DeviceOperator(A, tiled="auto")
	assert not op.tiled
	need = ctypes.c_size_t(1)
	_ffi.check(_ffi.lib().slq_dense_tiled_bytes(op._handle, ctypes.byref(need)))
	assert need.value == 0
	op64 = engine.DeviceOperator(np.eye(448), tiled="auto")
	_ffi.check(_ffi.lib().slq_dense_tiled_bytes(op64._handle, ctypes.byref(need)))
	assert op64.tiled and need.value % (16 * 8) == 0 and need.value >= 448 * 448 * 8


def test_lanczos_on_tiled_dense_operator_matches_reference(oracle):
	"""End to end through the Lanczos driver: alpha / beta of a 1000 x 1000 SPD operator, k = 32, vs the reference."""
	from primate_b200.operators import MatrixFunction
	from primate_b200.random import ProbeStream

	rng = np.random.default_rng(0)
	B = rng.standard_normal((1000, 1000))
	A = B @ B.T / 1000 + 0.1 * np.eye(1000)
	M = MatrixFunction(A, "log", deg=20, orth=0)
	assert M.device_op.tiled
	X = ProbeStream(1000, "rademacher", seed=5).block(32, 0)
	W = X.cpu().numpy().copy()
	_, out = M.quad_block(X, keep_tridiag=True)
	al, be = out["alpha"].cpu().numpy(), out["beta"].cpu().numpy()
	for p in (0, 17, 31):
		ra, rb, _ = oracle.ref_lanczos(A, W[:, p], 20, 1e-8, 0, ncv=20)
		assert np.max(np.abs(al[p, :20] - ra[:20]) / np.abs(ra[:20])) < 1e-10
		assert np.max(np.abs(be[p, 1:20] - rb[1:20]) / np.abs(rb[1:20])) < 1e-10
