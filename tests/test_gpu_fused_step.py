"""The fused Lanczos step (one kernel per step: k_stepA_tma<STEP_FUSED>, taken for CSR operators with rows of <= 64
non-zeros at orth = 0 when no basis is wanted) against the two-kernel step (step A + axpy, taken here by asking for the
basis) on the same start blocks.  The fused step forms v_{j+1} = w_j - alpha_j q_j on the fly in the NEXT step and scales
it with a norm computed from three fp64 sums (S_w - 2 cf <v, w> + cf^2 S_v); what it records as beta_j is the exact
norm of the formed vector, one step late, and the early break of lanczos.h:139-143 is decided on that exact value."""
import numpy as np
import pytest
import scipy.sparse as sp

from .conftest import laplacian_2d, rel_err

pytestmark = pytest.mark.gpu


def _both(A, X, deg, dtype=np.float64):
	"""(fused, two-kernel) Lanczos outputs for the host start block X (n, k)."""
	import torch

	from primate_b200 import engine

	op = engine.DeviceOperator(A, dtype=dtype, fused=True)
	Xd = torch.from_numpy(np.ascontiguousarray(X, dtype=dtype)).cuda()
	fused = op.lanczos(op._to_internal(Xd.clone()), deg, 1e-8, 0, 2)
	plain = op.lanczos(op._to_internal(Xd.clone()), deg, 1e-8, 0, max(deg, 2), want_basis=True)
	get = lambda o: (o["alpha"].cpu().numpy(), o["beta"].cpu().numpy(), o["steps"].cpu().numpy(), o["xnorm2"].cpu().numpy())  # noqa: E731
	return get(fused), get(plain)


@pytest.mark.parametrize("k", [8, 16, 32, 64])
@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-11), (np.float32, 2e-4)])
def test_fused_step_matches_two_kernel_step(k, dtype, tol):
	A = (laplacian_2d(48) + 0.3 * sp.identity(48 * 48)).astype(dtype)
	X = np.random.default_rng(k).standard_normal((A.shape[0], k))
	deg = 30
	(fa, fb, fs, fx), (pa, pb, ps, px) = _both(A, X, deg, dtype)
	assert np.array_equal(fs, ps) and np.all(fs == deg)
	assert np.array_equal(fx, px)
	assert rel_err(fa[:, :deg], pa[:, :deg]) < tol
	assert rel_err(fb[:, 1:deg], pb[:, 1:deg]) < tol  # exact norms of the formed vectors
	assert rel_err(fb[:, deg], pb[:, deg]) < max(tol, 1e-9)  # the last beta is the three-sum value (never used by the quadrature)


def test_fused_step_vs_reference_at_c2_parameters(oracle):
	"""C2's kernel shape (fp64, k = 64, deg 50) on a 200 x 200 grid: alpha / beta to the north-star's 1e-10 vs the reference."""
	import torch

	from primate_b200 import engine

	A = laplacian_2d(200)
	n = A.shape[0]
	st, inc = oracle.pcg64_state(7)
	X = engine.rademacher_block(st, inc, 0, 64, n, np.float64)
	W = X.cpu().numpy().copy()
	op = engine.DeviceOperator(A, fused=True)
	out = op.lanczos(op._to_internal(X), 50, 1e-8, 0, 2)
	al, be = out["alpha"].cpu().numpy(), out["beta"].cpu().numpy()
	for p in (0, 31, 63):
		ra, rb, _ = oracle.ref_lanczos(A, W[:, p], 50, 1e-8, 0, ncv=2)
		assert rel_err(al[p, :50], ra[:50]) < 1e-10 and rel_err(be[p, 1:50], rb[1:50]) < 1e-10
	assert torch.all(out["steps"] == 50)


@pytest.mark.parametrize("scale", [1.0, 1e4])
def test_fused_step_breakdown_is_decided_on_the_exact_norm(scale):
	"""Start vectors in a 3-dimensional invariant subspace: beta_3 is rounding noise.  The three-sum norm cannot resolve
	it (cancellation floor ~1e-8 ||w||, above the break tolerance once ||A|| is large), the exact norm one step later can:
	both paths stop after 3 steps, tails stay zero, beta_3 below the tolerance."""
	n = 4096
	d = np.repeat([1.0, 2.0, 5.0], [1000, 1000, 2096]) * scale
	A = sp.diags(d).tocsr()
	X = np.random.default_rng(3).standard_normal((n, 8))
	(fa, fb, fs, _), (pa, pb, ps, _) = _both(A, X, 12)
	assert np.all(fs == 3) and np.all(ps == 3)
	assert rel_err(fa[:, :3], pa[:, :3]) < 1e-10 and rel_err(fb[:, 1:3], pb[:, 1:3]) < 1e-10
	assert np.all(fa[:, 3:] == 0) and np.all(fb[:, 4:] == 0)
	assert np.all(fb[:, 3] < np.sqrt(n) * 1e-8 * max(scale, 1.0))


def test_fused_step_mixed_breakdown_columns_and_degree_one():
	"""Columns that break down next to columns that do not; deg = 1 (a single fused launch)."""
	n = 2500
	A = (laplacian_2d(50) + sp.identity(n)).tocsr()
	rng = np.random.default_rng(5)
	X = rng.standard_normal((n, 8))
	ew, ev = np.linalg.eigh(A.toarray())
	X[:, 2] = ev[:, 10]  # an eigenvector: breaks down after one step
	X[:, 5] = ev[:, 3] + 2.0 * ev[:, 700]  # two-dimensional invariant subspace
	(fa, fb, fs, _), (pa, pb, ps, _) = _both(A, X, 20)
	assert np.array_equal(fs, ps) and fs[2] == 1 and fs[5] == 2 and np.all(np.delete(fs, [2, 5]) == 20)
	for p in range(8):
		m = int(fs[p])
		assert rel_err(fa[p, :m], pa[p, :m]) < 1e-9 and (m < 2 or rel_err(fb[p, 1:m], pb[p, 1:m]) < 1e-9)
	assert abs(fa[2, 0] - ew[10]) < 1e-12 * ew[10]
	(fa1, fb1, fs1, _), (pa1, pb1, ps1, _) = _both(A, X, 1)
	assert np.all(fs1 == 1) and rel_err(fa1[:, 0], pa1[:, 0]) < 1e-12


def test_fused_step_policy_and_switch():
	"""`fused="auto"`: on for a mesh in the locality order (C2's operator: 0.74 off-tile non-zeros per row), off for a 3-D
	7-point Laplacian (2.7 per row even after the ordering: the doubled gathers cost more than the saved panel passes);
	`fused=False` takes the two-kernel step (same results as asking for the basis)."""
	import torch

	from primate_b200 import engine

	A2 = laplacian_2d(300)  # n = 90 000 >= the reorder threshold
	op2 = engine.DeviceOperator(A2)
	assert op2.perm is not None and op2.offtile_per_row < 1.2 and op2.fused_step
	T = sp.diags([-1.0, 2.0, -1.0], [-1, 0, 1], shape=(44, 44))
	I = sp.identity(44)  # noqa: E741
	A3 = (sp.kron(sp.kron(T, I), I) + sp.kron(sp.kron(I, T), I) + sp.kron(sp.kron(I, I), T)).tocsr()
	op3 = engine.DeviceOperator(A3)
	assert op3.offtile_per_row > 2.0 and not op3.fused_step
	X = np.random.default_rng(0).standard_normal((A2.shape[0], 16))
	Xd = torch.from_numpy(X).cuda()
	outs = []
	for fused in (True, False):
		op = engine.DeviceOperator(A2, fused=fused)
		assert op.fused_step == fused
		o = op.lanczos(op._to_internal(Xd.clone()), 25, 1e-8, 0, 2)
		outs.append((o["alpha"].cpu().numpy(), o["beta"].cpu().numpy()))
	assert rel_err(outs[0][0][:, :25], outs[1][0][:, :25]) < 1e-11 and rel_err(outs[0][1][:, 1:25], outs[1][1][:, 1:25]) < 1e-11
	assert not np.array_equal(outs[0][1][:, 25], np.zeros(16))


@pytest.mark.parametrize("fused", [True, False])
def test_lanczos_does_not_read_unwritten_workspace(fused):
	"""The caller's workspace arrives with arbitrary contents: poisoned with NaN bit patterns between two identical calls,
	the coefficients must not change (the first fused step multiplies beta_0 = 0 by the scale of a "previous" vector)."""
	import torch

	from primate_b200 import engine

	A = (laplacian_2d(40) + 0.5 * sp.identity(1600)).tocsr()
	X = torch.from_numpy(np.random.default_rng(2).standard_normal((1600, 16))).cuda()
	op = engine.DeviceOperator(A, fused=fused)
	a = op.lanczos(X.clone(), 12, 1e-8, 0, 2)
	for buf in op._ws.values():
		buf.fill_(0xFF)  # every double / float in the workspace becomes a NaN
	b = op.lanczos(X.clone(), 12, 1e-8, 0, 2)
	for key in ("alpha", "beta", "steps", "xnorm2"):
		assert torch.equal(a[key], b[key]), key
	assert torch.isfinite(b["alpha"]).all() and torch.isfinite(b["beta"]).all()
