"""CPU tests of the host-side Monte-Carlo driver (primate_b200.trace) and of probe sharding over ranks
(primate_b200.sharding; gloo, world_size 2), with the quadrature samples injected from the oracle / the golden
vectors of the unmodified reference -- no device compute here."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _table_sampler(samples):
	calls = []

	def fn(first, k):
		calls.append((first, k))
		assert k & (k - 1) == 0 and 1 <= k <= 64  # device blocks are powers of two
		return np.asarray(samples[first : first + k], dtype=np.float64)

	fn.calls = calls
	return fn


def test_fixed_count_and_batched_match_reference_golden(golden):
	from primate_b200 import sharding, trace
	from primate_b200.estimators import CountCriterion, EstimatorResult, MeanEstimator

	S = golden["hutch_log_count200_seed5_full_b32_samples"]
	solo = sharding.SoloComm()
	est = MeanEstimator(covariance=True)
	used = trace._run_fixed_count(_table_sampler(S), est, 200, solo, 32)
	assert used == 200 and len(est) == 200
	assert abs(est.estimate - golden["hutch_log_count200_seed5"].item()) < 1e-10
	# full=True cadence: 7 batches of 32 = 224 samples (trace.py:107-110)
	est = MeanEstimator(covariance=True)
	cc = CountCriterion(200)
	seen = []
	used = trace._run_batched(_table_sampler(S), est, cc, 32, lambda r: seen.append(len(r.estimator)), EstimatorResult(est, cc), solo, 32)
	assert used == 224 and seen == [32 * i for i in range(1, 8)]
	assert abs(est.estimate - golden["hutch_log_count200_seed5_full_b32"].item()) < 1e-10


def test_default_criterion_replay_stops_where_the_reference_stops(oracle, golden):
	"""Count(200) | Confidence(.95, atol=1): per-probe cadence replayed on blocks of 32 speculative probes."""
	from primate_b200 import sharding, trace
	from primate_b200.estimators import ConfidenceCriterion, CountCriterion, MeanEstimator

	A = golden["A150"]
	n_ref = int(golden["hutch_log_default_seed11_n"])
	W = oracle.ref_isotropic((150, 256), "rademacher", seed=11)
	S = oracle.ref_quad(A, W, "log", deg=20, orth=0)
	est = MeanEstimator(covariance=True)
	cc = CountCriterion(200) | ConfidenceCriterion(0.95, atol=1.0, rtol=0.0)
	used = trace._run_per_probe(_table_sampler(S), est, cc, sharding.SoloComm(), 32)
	assert used == n_ref == len(est)
	assert abs(est.estimate - golden["hutch_log_default_seed11"].item()) < 1e-9
	# an adaptive stop in the middle of a block
	n8, e8 = int(golden["hutch_log_conf_seed3_b8_n"]), golden["hutch_log_conf_seed3_b8"].item()
	W = oracle.ref_isotropic((150, n8 + 8), "rademacher", seed=3)
	S = oracle.ref_quad(A, W, "log", deg=20, orth=0)
	est = MeanEstimator(covariance=True)
	cc = ConfidenceCriterion(0.9, atol=2.0, rtol=0.0)
	from primate_b200.estimators import EstimatorResult

	used = trace._run_batched(_table_sampler(S), est, cc, 8, lambda r: None, EstimatorResult(est, cc), sharding.SoloComm(), 32)
	assert used == n8 and abs(est.estimate - e8) < 1e-9


def test_probe_range_and_merge_records():
	from primate_b200 import sharding

	for total, size in ((10000, 8), (7, 3), (5, 8), (0, 2)):
		got = [sharding.probe_range(total, r, size) for r in range(size)]
		flat = [i for lo, hi in got for i in range(lo, hi)]
		assert flat == list(range(total))
	rng = np.random.default_rng(0)
	x = rng.normal(1e6, 1e3, size=1000)
	parts = np.array_split(x, 7)
	n, mu, S = sharding.merge_records([sharding.record_of(p) for p in parts])
	assert n == 1000 and abs(mu - x.mean()) < 1e-9 * abs(x.mean())
	assert abs(S / (n - 1) - x.var(ddof=1)) < 1e-10 * x.var(ddof=1)


# ---- world_size-2 gloo: the sharded driver gives every rank the single-process result ---------------------------
_WORKER = r"""
import os, sys, json
import numpy as np
sys.path.insert(0, sys.argv[1])
import torch.distributed as dist
dist.init_process_group("gloo")
from primate_b200 import sharding, trace
from primate_b200.estimators import ConfidenceCriterion, CountCriterion, EstimatorResult, MeanEstimator
S = np.load(os.path.join(sys.argv[1], "tests", "golden", "slq_golden.npz"))["hutch_log_count200_seed5_full_b32_samples"]
S = np.concatenate([S, S[::-1] * 1.01, S * 0.99])  # 672 deterministic "samples"
comm = sharding.resolve_comm("auto")
assert comm.size == 2
seen = []
def sampler(first, k):
    seen.append((first, k))
    return np.asarray(S[first:first + k])
out = {}
# (1) fixed count, records path (count, mean, M2 per rank; one all-gather)
est = MeanEstimator(covariance=True)
stats = lambda y: np.array(sharding.record_of(y))
trace._run_fixed_count(sampler, est, 200, comm, 32, stats)
out["fixed_records"] = [len(est), est.estimate, float(est._cov.covariance())]
out["fixed_owned"] = [min(f for f, _ in seen), max(f + k for f, k in seen)]
# (2) fixed count, samples path (record=True)
est = MeanEstimator(covariance=True, record=True)
trace._run_fixed_count(sampler, est, 201, comm, 32, stats)
out["fixed_samples"] = [len(est), est.estimate, float(np.ravel(est.values)[-1])]
# (3) adaptive per-probe
est = MeanEstimator(covariance=True)
cc = CountCriterion(300) | ConfidenceCriterion(0.95, atol=1.2, rtol=0.0)
used = trace._run_per_probe(sampler, est, cc, comm, 32)
out["adaptive"] = [used, len(est), est.estimate]
# (4) batched cadence
est = MeanEstimator(covariance=True)
cc = CountCriterion(100)
used = trace._run_batched(sampler, est, cc, 24, lambda r: None, EstimatorResult(est, cc), comm, 32)
out["batched"] = [used, len(est), est.estimate]
# (5) block drivers: columns of a block product split over the ranks, panels all-gathered
rng = np.random.default_rng(0)
B = rng.standard_normal((40, 40)); B = B + B.T
X = rng.standard_normal((40, 7))
Y = sharding.sharded_columns(lambda Z: B @ Z, X, comm, chunk_bytes=256)  # several pieces
q = sharding.sharded_columns(lambda Z: np.einsum("ij,ij->j", Z, B @ Z)[None, :], X, comm).ravel()
y1 = sharding.sharded_columns(lambda Z: B @ Z, X[:, :1], comm)  # a single column is not split
out["block"] = [float(np.abs(Y - B @ X).max()), float(np.abs(q - np.einsum("ij,ij->j", X, B @ X)).max()),
                float(np.abs(y1 - B @ X[:, :1]).max()), float(Y.sum()), float(q.sum())]
print("RESULT " + json.dumps(out), flush=True)
dist.destroy_process_group()
"""


def _free_port():
	with socket.socket() as s:
		s.bind(("127.0.0.1", 0))
		return s.getsockname()[1]


def test_gloo_world2_sharded_driver_matches_single_process(golden, tmp_path):
	import json
	import subprocess

	from primate_b200 import sharding, trace
	from primate_b200.estimators import ConfidenceCriterion, CountCriterion, EstimatorResult, MeanEstimator

	script = tmp_path / "worker.py"
	script.write_text(_WORKER)
	port = _free_port()
	procs = []
	for r in range(2):
		env = dict(os.environ, RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
		           CUDA_VISIBLE_DEVICES="")  # fmt: skip
		procs.append(subprocess.Popen([sys.executable, str(script), ROOT], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
	outs = []
	for p in procs:
		o, e = p.communicate(timeout=240)
		assert p.returncode == 0, e[-2000:]
		outs.append(json.loads([ln for ln in o.splitlines() if ln.startswith("RESULT ")][-1][7:]))
	assert outs[0]["fixed_records"] == outs[1]["fixed_records"] and outs[0]["adaptive"] == outs[1]["adaptive"]
	assert outs[0]["fixed_owned"] == [0, 100] and outs[1]["fixed_owned"][0] == 100
	assert max(outs[0]["block"][:3]) < 1e-12 and outs[0]["block"] == outs[1]["block"]  # identical bytes on both ranks

	S = golden["hutch_log_count200_seed5_full_b32_samples"]
	S = np.concatenate([S, S[::-1] * 1.01, S * 0.99])
	solo = sharding.SoloComm()
	tab = _table_sampler(S)
	est = MeanEstimator(covariance=True)
	trace._run_fixed_count(tab, est, 200, solo, 32)
	n, mu, var = outs[0]["fixed_records"]
	assert n == 200 and abs(mu - est.estimate) < 1e-9 * abs(mu) and abs(var - est._cov.covariance()) < 1e-9 * var
	assert abs(mu - np.mean(S[:200])) < 1e-9 * abs(mu)
	n, mu, last = outs[0]["fixed_samples"]
	assert n == 201 and abs(mu - np.mean(S[:201])) < 1e-9 * abs(mu) and last == S[200]
	est = MeanEstimator(covariance=True)
	cc = CountCriterion(300) | ConfidenceCriterion(0.95, atol=1.2, rtol=0.0)
	used = trace._run_per_probe(tab, est, cc, solo, 32)
	assert outs[0]["adaptive"][0] == used and abs(outs[0]["adaptive"][2] - est.estimate) < 1e-12 * abs(est.estimate)
	est = MeanEstimator(covariance=True)
	cc = CountCriterion(100)
	used = trace._run_batched(tab, est, cc, 24, lambda r: None, EstimatorResult(est, cc), solo, 32)
	assert outs[0]["batched"][0] == used == 120 and abs(outs[0]["batched"][2] - est.estimate) < 1e-12 * abs(est.estimate)


def test_c_abi_exports_every_declared_symbol():
	"""libslq.so loads on a CPU-only box and exports each function include/slq.h declares (no compute calls)."""
	import re

	from primate_b200 import _ffi

	with open(os.path.join(ROOT, "include", "slq.h")) as f:
		hdr = f.read()
	hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
	declared = set(re.findall(r"\b(slq_[a-z0-9_]+)\s*\(", hdr)) - {"slq_matmat_fn"}
	L = _ffi.lib()
	for name in declared:
		assert hasattr(L, name), f"{name} declared in slq.h but not exported"
	assert declared == set(_ffi.EXPORTED), declared ^ set(_ffi.EXPORTED)
	assert L.slq_version() == 100


def test_product_never_imports_the_oracle():
	for dirpath, _, files in os.walk(os.path.join(ROOT, "primate_b200")):
		for fn in files:
			if fn.endswith((".py", ".cu", ".cuh", ".h")):
				src = open(os.path.join(dirpath, fn)).read()
				assert "oracle" not in src.replace("no oracle", ""), f"{fn} mentions the oracle"


def test_hutchpp_xtrace_xdiag_diag_host_algebra_match_reference_golden(golden):
	"""Block drivers on a plain ndarray (NumPy products): the leave-one-out / deflation algebra equals the reference's."""
	from primate_b200.diagonal import diag, xdiag
	from primate_b200.trace import hutchpp, xtrace

	A = golden["A150"]
	assert abs(hutchpp(A, m=30, seed=4) - golden["hutchpp_A_m30_seed4"].item()) < 1e-9
	assert np.allclose(xdiag(A, m=20, seed=6), golden["xdiag_A_m20_seed6"], rtol=1e-9, atol=1e-11)
	assert abs(xtrace(A, batch=10, seed=8) - golden["xtrace_A_b10_seed8"].item()) < 1e-8
	assert np.allclose(diag(A, converge="count", count=50, seed=10), golden["diag_A_seed10"], rtol=1e-10, atol=1e-12)


def test_no_library_linear_algebra_on_device_panels():
	"""The device-panel drivers own their algebra: no torch.linalg / torch matmul on n x m panels (CholeskyQR2 and the
	panel kernels of libslq instead); NumPy's dense solvers only ever see m x m matrices."""
	import re

	for fn in ("trace.py", "diagonal.py", "panels.py", "operators.py", "engine.py", "random.py"):
		src = open(os.path.join(ROOT, "primate_b200", fn)).read()
		assert "torch.linalg" not in src and "linalg.qr(" not in src.replace("np.linalg.qr(", ""), fn
		assert not re.search(r"\btorch\.(matmul|mm|bmm)\b", src), fn


def test_update_trinv_and_block_append():
	"""src/primate/linalg.py:5-24 (tests/test_linalg.py:5-27): appending columns to an upper-triangular inverse."""
	from primate_b200.linalg import append_trinv, update_trinv

	rng = np.random.default_rng(0)
	R = np.triu(rng.standard_normal((9, 9))) + 4 * np.eye(9)
	Rinv = np.zeros((0, 0))
	for j in range(9):
		Rinv = update_trinv(Rinv, R[: j + 1, j])
		assert np.allclose(Rinv, np.linalg.inv(R[: j + 1, : j + 1]), atol=1e-10)
	blk = append_trinv(np.linalg.inv(R[:5, :5]), R[:5, 5:], np.linalg.inv(R[5:, 5:]))
	assert np.allclose(blk, np.linalg.inv(R), atol=1e-10)


def test_gathered_samples_keep_genuine_nans():
	"""Multi-rank sample tables are cut by each rank's known share, not by filtering NaN padding: a NaN quadrature sample
	(an inv / sqrt of a bad Ritz value) reaches the estimator as it would on one rank."""
	from primate_b200 import trace

	class FakeComm:
		rank, size, backend = 1, 3, "fake"

		def __init__(self, rows):
			self.rows = rows

		def all_gather(self, x):
			return np.stack(self.rows)

	rows = [np.array([1.0, np.nan, 3.0]), np.array([4.0, 5.0, 6.0]), np.array([7.0, 0.0, 0.0])]  # 7 probes over 3 ranks: 3 + 3 + 1
	got = trace._gather_ranges(rows[1], 7, FakeComm(rows))
	assert got.shape == (7,) and np.isnan(got[1]) and np.array_equal(got[[0, 2, 3, 4, 5, 6]], [1.0, 3.0, 4.0, 5.0, 6.0, 7.0])


def test_locality_ordering_is_remembered_per_sparsity_structure():
	"""engine._cluster_order: the second request for the same structure returns the remembered permutation without
	calling the clustering again; another structure (or tile) is clustered afresh; every result is a permutation."""
	import scipy.sparse as sp

	from primate_b200 import _ffi, engine

	from .conftest import laplacian_2d

	L = _ffi.lib()
	engine._ORDER_CACHE.clear()
	A = laplacian_2d(40)
	ip, ix = np.ascontiguousarray(A.indptr, dtype=np.int32), np.ascontiguousarray(A.indices, dtype=np.int32)
	p1 = engine._cluster_order(L, A.shape[0], ip, ix, 120)
	assert len(engine._ORDER_CACHE) == 1 and np.array_equal(np.sort(p1), np.arange(A.shape[0]))
	p2 = engine._cluster_order(L, A.shape[0], ip.copy(), ix.copy(), 120)
	assert p2 is p1  # same structure in other arrays: remembered
	p3 = engine._cluster_order(L, A.shape[0], ip, ix, 60)
	assert p3 is not p1 and len(engine._ORDER_CACHE) == 2
	B = (A + sp.diags([1.0], [7], shape=A.shape) + sp.diags([1.0], [-7], shape=A.shape)).tocsr()
	bp, bx = np.ascontiguousarray(B.indptr, dtype=np.int32), np.ascontiguousarray(B.indices, dtype=np.int32)
	p4 = engine._cluster_order(L, B.shape[0], bp, bx, 120)
	assert p4 is not p1 and np.array_equal(np.sort(p4), np.arange(B.shape[0]))
	for i in range(5):  # bounded
		engine._cluster_order(L, A.shape[0], ip, ix, 100 + i)
	assert len(engine._ORDER_CACHE) <= 4


def test_device_side_permutation_equals_slq_csr_permute():
	"""engine._permute_csr_device (torch index bookkeeping; CPU tensors here) produces exactly the arrays of the host
	routine slq_csr_permute -- rows in the new order, columns relabelled and sorted -- so moving the permutation to the
	device changed no operator bit."""
	import scipy.sparse as sp
	import torch

	from primate_b200 import _ffi, engine

	from .conftest import laplacian_2d

	L = _ffi.lib()
	rng = np.random.default_rng(0)
	R = sp.random(2000, 2000, density=0.004, random_state=1)
	R = (R + R.T).tolil()
	R[5, :] = 1.0
	R[:, 5] = 1.0  # one long row / column; some rows stay empty
	for A, dtype, code in ((laplacian_2d(50), np.float64, _ffi.SLQ_F64), (R.tocsr(), np.float32, _ffi.SLQ_F32)):
		C = A.tocsr().astype(dtype)
		C.sort_indices()
		n = C.shape[0]
		ip, ix = np.ascontiguousarray(C.indptr, dtype=np.int32), np.ascontiguousarray(C.indices, dtype=np.int32)
		va = np.ascontiguousarray(C.data, dtype=dtype)
		for perm in (engine._cluster_order(L, n, ip, ix, 120), rng.permutation(n).astype(np.int32)):
			ip2, ix2, va2 = np.empty_like(ip), np.empty_like(ix), np.empty_like(va)
			_ffi.check(L.slq_csr_permute(n, ip.ctypes.data, ix.ctypes.data, va.ctypes.data, code, perm.ctypes.data, ip2.ctypes.data,
			                             ix2.ctypes.data, va2.ctypes.data))  # fmt: skip
			r, c, v = engine._permute_csr_device(torch.from_numpy(ip), torch.from_numpy(ix), torch.from_numpy(va), torch.from_numpy(perm), n)
			assert np.array_equal(r.numpy(), ip2) and np.array_equal(c.numpy(), ix2) and np.array_equal(v.numpy(), va2)


def test_offtile_estimate_matches_brute_force(monkeypatch):
	"""engine._offtile_per_row (sampled tiles, host arrays) and its torch twin against a brute-force count; the numbers
	behind the fused-step policy: a 2-D mesh in natural order ~2 off-tile non-zeros per row, ~0.75 after the ordering."""
	import torch

	from primate_b200 import _ffi, engine

	from .conftest import laplacian_2d

	A = laplacian_2d(150)
	n = A.shape[0]
	ip, ix = np.ascontiguousarray(A.indptr, dtype=np.int32), np.ascontiguousarray(A.indices, dtype=np.int32)
	rows = np.repeat(np.arange(n), np.diff(ip))
	brute = np.count_nonzero(rows // 120 != ix // 120) / n
	est = engine._offtile_per_row(ip, ix, sample=10**6)  # every tile
	assert abs(est - brute) < 1e-12
	assert abs(engine._offtile_per_row(ip, ix) - brute) < 0.1  # the default sample
	monkeypatch.setattr(engine, "_torch", lambda: torch)
	dev = engine._offtile_per_row_device(torch.from_numpy(ip), torch.from_numpy(ix), n, sample=10**6)
	assert abs(dev - brute) < 1e-12
	perm = engine._cluster_order(_ffi.lib(), n, ip, ix, 120)
	r, c, _ = engine._permute_csr_device(torch.from_numpy(ip), torch.from_numpy(ix), torch.from_numpy(A.data.copy()), torch.from_numpy(perm), n)
	after = engine._offtile_per_row(r.numpy(), c.numpy())
	assert brute > 1.5 and after < 1.0 and after <= engine.FUSED_MAX_OFFTILE < brute
