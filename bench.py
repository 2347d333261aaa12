"""bench.py -- Lanczos probe-steps/s of hutch log-det (BASELINE.json metric) on config C2.

  python bench.py [--gpus N] [--steps K] [--warmup W]            # our arm (libslq on B200)
  python bench.py --impl reference [--gpus N] [--steps K] ...     # the reference's own CPU code on the host cores

Workload (BASELINE.json configs[1], SURVEY 8d "C2"): 5-point Laplacian on a 1000x1000 grid (n = 1e6, CSR fp64 /
int32, 63.95 MB), f = log, deg = 50, orth = 0, Rademacher probes.  One STEP = one hutch pass over the WHOLE job of
`--probes` (default 10 000) probes, sharded over the N ranks -- STRONG scaling, the north-star's ">= 7x at 8 GPUs" is
about this job: in step s rank r owns the probe indices s*P + probe_range(P, r, N) of ONE PCG64 stream (device
jump-ahead, no communication) and the only collective is one all-gather of each rank's (count, mean, M2) record per
step.  1 probe-step = one Lanczos iteration of one probe (lanczos.h:123-148).  For N > 1 a `weak` sub-dict carries the
round-1 reading as well (P probes PER RANK per step).

  value   device-resident throughput: operator and probe generator state already in HBM, timed with CUDA events;
  e2e     the same metric through the public API with HOST buffers: every step builds a MatrixFunction from the
          host SciPy CSR matrix (pinned arrays -> H2D upload inside the timed region, on every rank), runs
          hutch(M, converge="count", count=P) and reads the estimate back;
  roofline   dominant kernel (fused CSR SpMM + three-term recurrence, "step A"): algorithmic bytes per launch /
          its average duration measured with CUDA events on the launching stream inside the timed region;
  cpu_baseline  the reference's compiled extension (oracle/_ref) driven exactly as hutch drives it, on a bounded
          sample of the same workload on the host cores (N = 1 only);
  configs  (N = 1) the other BASELINE configs at stated sizes, each with its throughput and roofline fraction.

Working set per block (3 panels of n x 64 fp64 = 1.5 GB + CSR 64 MB) is far larger than the 126 MB L2, so no L2
flush is needed between timed iterations ("inputs larger than L2").
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
	sys.path.insert(0, ROOT)

METRIC = "lanczos_probe_steps_per_sec_hutch_logdet"
UNIT = "probe-steps/s"


def parse_args():
	ap = argparse.ArgumentParser()
	ap.add_argument("--gpus", type=int, default=1)
	ap.add_argument("--steps", type=int, default=3)
	ap.add_argument("--warmup", type=int, default=3)
	ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
	ap.add_argument("--grid", type=int, default=1000, help="Laplacian grid side (C2: 1000 -> n = 1e6)")
	ap.add_argument("--deg", type=int, default=50)
	ap.add_argument("--probes", type=int, default=10000, help="probes per step, the whole job (C2: 10 000); sharded over the ranks")
	ap.add_argument("--block", type=int, default=64, help="probes per device block (power of two <= 64)")
	ap.add_argument("--cpu-probes", type=int, default=0, help="reference arm: probes per worker per step (0 = auto)")
	ap.add_argument("--cpu-workers", type=int, default=0, help="reference arm: worker processes (0 = all host cores)")
	ap.add_argument("--no-cpu-baseline", action="store_true")
	ap.add_argument("--no-e2e", action="store_true")
	ap.add_argument("--no-configs", action="store_true", help="skip the C1/C3/C4/C5 side measurements (N = 1)")
	ap.add_argument("--no-weak", action="store_true", help="N > 1: skip the secondary weak-scaling pass")
	ap.add_argument("--c3-scale", type=int, default=22, help="configs: R-MAT scale of the C3 side measurement (BASELINE: 24)")
	return ap.parse_args()


def laplacian_2d(g: int):
	import scipy.sparse as sp

	T = sp.diags([-1.0, 2.0, -1.0], [-1, 0, 1], shape=(g, g))
	A = (sp.kron(sp.identity(g), T) + sp.kron(T, sp.identity(g))).tocsr().astype(np.float64)
	A.sort_indices()
	return A


def workload_name(args) -> str:
	return (f"C2: hutch logdet, 2-D 5-point Laplacian {args.grid}x{args.grid} (n={args.grid ** 2} CSR fp64/int32), "
	        f"f=log, deg={args.deg}, orth=0, {args.probes} Rademacher probes per step (the whole job, sharded over the ranks)")  # fmt: skip


def config_of(args) -> dict:
	"""The `config` object, identical in both arms (everything arm-specific goes to `details`)."""
	return {"workload": workload_name(args), "grid": args.grid, "deg": args.deg, "probes_per_step": args.probes, "fun": "log"}


# ---- clocks -----------------------------------------------------------------------------------------------------
class ClockSampler:
	"""Samples nvidia-smi clocks / throttle reasons of one GPU every 200 ms while the timed region runs."""

	Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
	     "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")  # fmt: skip

	def __init__(self, index: int = 0):
		self.index, self.proc, self.lines = index, None, []

	def start(self):
		try:
			self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
			                              "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)  # fmt: skip
			self.thread = threading.Thread(target=self._pump, daemon=True)
			self.thread.start()
		except OSError:
			self.proc = None

	def _pump(self):
		for line in self.proc.stdout:
			self.lines.append(line.strip())

	def stop(self) -> dict:
		if self.proc is None:
			return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
		self.proc.terminate()
		try:
			self.proc.wait(timeout=5)
		except subprocess.TimeoutExpired:
			self.proc.kill()
		sm, smax, power, reasons = [], [], [], set()
		names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
		for ln in self.lines:
			f = [x.strip() for x in ln.split(",")]
			if len(f) < 7:
				continue
			try:
				sm.append(float(f[0]))
				smax.append(float(f[1]))
				power.append(float(f[2]))
			except ValueError:
				continue
			for nm, val in zip(names, f[3:7]):
				if val.lower().startswith("active"):
					reasons.add(nm)
		if not sm:
			return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
		busy = [c for c, p in zip(sm, power) if p > 0.5 * max(power)] or sm
		return {"sm_mhz": float(np.median(busy)), "sm_max_mhz": float(max(smax)), "reasons": sorted(reasons),
		        "samples": len(sm), "power_w_max": float(max(power))}  # fmt: skip


# ---- reference arm: the reference's own CPU implementation on the host cores -----------------------------------------
_W = {}


def _ref_worker_init(grid: int, deg: int):
	os.environ.setdefault("OMP_NUM_THREADS", "1")
	from oracle import slq_oracle

	_W["A"] = laplacian_2d(grid)
	_W["deg"] = deg
	_W["orc"] = slq_oracle
	slq_oracle.reference_ext()  # oracle/_ref: the reference's compiled `_lanczos` extension


def _ref_worker_run(task):
	first, count = task
	orc = _W["orc"]
	t0 = time.perf_counter()
	_, samples = orc.ref_hutch(_W["A"], fun="log", deg=_W["deg"], orth=0, count=count, seed=0, batch=1, first_probe=first,
	                           return_samples=True)  # fmt: skip
	return samples, time.perf_counter() - t0


def host_cores() -> int:
	try:
		return len(os.sched_getaffinity(0))
	except AttributeError:
		return os.cpu_count() or 1


def run_reference(args, emit: bool = True) -> dict:
	"""Times oracle/_ref (reference `_lanczos` extension + the hutch glue of trace.py:107-116 / operators.py:126-151)
	on a bounded sample: every step, each of W worker processes runs `p` probes of the C2 job (disjoint probe indices of
	the one PCG64 stream, PCG64.advance: SURVEY 8d).  The reference is single-threaded per probe, so
	the host cores are used by sharding probes over processes."""
	import multiprocessing as mp

	rank = int(os.environ.get("RANK", "0"))
	if rank != 0:
		return {}
	workers = args.cpu_workers or host_cores()
	workers = max(1, min(workers, 256))
	per = args.cpu_probes or 2
	ctx = mp.get_context("spawn")
	with ctx.Pool(workers, initializer=_ref_worker_init, initargs=(args.grid, args.deg)) as pool:
		def one_step(s: int):
			tasks = [((s * workers + w) * per, per) for w in range(workers)]
			t0 = time.perf_counter()
			out = pool.map(_ref_worker_run, tasks, chunksize=1)
			dt = time.perf_counter() - t0
			return dt, out

		for s in range(args.warmup):
			one_step(s)
		t_total, core_seconds, last = 0.0, 0.0, None
		for s in range(args.steps):
			dt, out = one_step(args.warmup + s)
			t_total += dt
			core_seconds += sum(o[1] for o in out)
			last = out
	probe_steps = args.steps * workers * per * args.deg
	value = probe_steps / t_total
	sample = (f"{workers} worker processes x {per} probes per step (disjoint probe ranges of the C2 stream, deg={args.deg}), "
	          f"{args.steps} steps; single-core rate {probe_steps / core_seconds:.1f} probe-steps/s/core")  # fmt: skip
	line = {
		"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
		"warmup": args.warmup, "ms_per_step": 1e3 * t_total / args.steps, "higher_is_better": True, "scaling": "strong",
		"vs_baseline": None, "dtype": "f64", "data": "synthetic",
		"config": config_of(args), "details": {"timing": "host wall clock (CPU only)"},
		"cpu_baseline": {"value": value, "unit": UNIT, "cores": workers, "kind": "reference", "sample": sample},
		"e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
		"gpu_launches": 0,
		"check": {"mean_logdet_sample": float(np.mean(np.concatenate([o[0] for o in last])))},
	}  # fmt: skip
	if emit:
		print(json.dumps(line), flush=True)
	return line


# ---- our arm -------------------------------------------------------------------------------------------------------
def pinned_csr(A):
	"""Re-homes the CSR arrays of `A` in pinned host memory (so the e2e upload is a pinned H2D copy)."""
	import scipy.sparse as sp
	import torch

	def pin(x, dt):
		t = torch.from_numpy(np.ascontiguousarray(x, dtype=dt)).pin_memory()
		return t, t.numpy()

	keep, arrs = [], []
	for x, dt in ((A.data, np.float64), (A.indices, np.int32), (A.indptr, np.int32)):
		t, a = pin(x, dt)
		keep.append(t)
		arrs.append(a)
	P = sp.csr_matrix((arrs[0], arrs[1], arrs[2]), shape=A.shape, copy=False)
	P.has_sorted_indices = True
	P._pinned_keepalive = keep
	return P


def run_ours(args) -> None:
	import torch

	rank = int(os.environ.get("RANK", "0"))
	local = int(os.environ.get("LOCAL_RANK", "0"))
	world = int(os.environ.get("WORLD_SIZE", "1"))
	if not torch.cuda.is_available():
		raise SystemExit("bench.py (our arm) needs a CUDA device: libslq has no CPU fallback")
	torch.cuda.set_device(local)
	import torch.distributed as dist

	if world > 1:
		os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
		if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
			os.environ["NCCL_DEBUG"] = "WARN"  # keep NCCL's version banner off stdout: rank 0 prints ONE JSON line
		dist.init_process_group("nccl", device_id=torch.device("cuda", local))
	from primate_b200 import _ffi, engine, sharding
	from primate_b200.operators import MatrixFunction
	from primate_b200.random import ProbeStream
	from primate_b200.trace import hutch

	L = _ffi.lib()
	comm = sharding.resolve_comm("auto")
	A = laplacian_2d(args.grid)
	n, nnz = A.shape[0], A.nnz
	P, deg, kb = args.probes, args.deg, args.block
	M = MatrixFunction(A, "log", deg=deg, orth=0)
	_ = M.device_op  # ingest once (locality ordering + upload): `value` is measured with the operator resident in HBM
	stream = ProbeStream(n, "rademacher", seed=0, dtype=np.float64)

	lo, hi = sharding.probe_range(P, rank, world)  # this rank's share of the P-probe job (strong scaling)

	def device_step(s: int, first=None, count=None):
		"""One hutch pass over this rank's share of the step's probes, everything resident; merged (count, mean, M2)."""
		first = s * P + lo if first is None else first
		count = hi - lo if count is None else count
		state = torch.zeros(3, dtype=torch.float64, device="cuda")
		if count > 0:
			y = M.quad_range(stream, first, count, kb)  # Lanczos blocks back to back, then ONE batched Golub-Welsch launch
			engine.stats_update(y, state)
		recs = comm.all_gather(state)  # one tiny collective per convergence check (N > 1: NCCL over NVLink)
		return sharding.merge_records([tuple(r) for r in recs])

	def sync_all():
		torch.cuda.synchronize()
		if world > 1:
			dist.barrier()
			torch.cuda.synchronize()

	for s in range(args.warmup):
		device_step(s)
	sync_all()
	clocks = ClockSampler(local)
	if rank == 0:
		clocks.start()
	L.slq_profile_enable(8192)
	launches0 = L.slq_launch_count()
	e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
	e0.record()
	rec = None
	for s in range(args.steps):
		rec = device_step(args.warmup + s)
	e1.record()
	sync_all()
	ms = e0.elapsed_time(e1)
	launches = L.slq_launch_count() - launches0
	clk = clocks.stop() if rank == 0 else None
	import ctypes

	prof = {}
	for cls, name in ((0, "stepA"), (1, "axpy")):
		tot, cnt, seen = ctypes.c_double(), ctypes.c_longlong(), ctypes.c_longlong()
		_ffi.check(L.slq_profile_read(cls, ctypes.byref(tot), ctypes.byref(cnt), ctypes.byref(seen)))
		prof[name] = (tot.value, cnt.value, seen.value)
	L.slq_profile_enable(0)
	t = torch.tensor([ms], dtype=torch.float64, device="cuda")
	if world > 1:
		dist.all_reduce(t, op=dist.ReduceOp.MAX)
	ms = float(t.item())
	total_probe_steps = P * deg * args.steps
	value = total_probe_steps / (ms * 1e-3)

	# ---- secondary: weak scaling (P probes PER RANK per step, the round-1 reading) ----------------------------------
	weak = None
	if world > 1 and not args.no_weak:
		wsteps = max(1, min(args.steps, 2))
		device_step(0, first=rank * P, count=P)
		sync_all()
		w0, w1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
		w0.record()
		for s_ in range(wsteps):
			device_step(0, first=((1 + s_) * world + rank) * P, count=P)
		w1.record()
		sync_all()
		wt = torch.tensor([w0.elapsed_time(w1)], dtype=torch.float64, device="cuda")
		dist.all_reduce(wt, op=dist.ReduceOp.MAX)
		weak = {"scaling": "weak", "probes_per_rank_per_step": P, "steps": wsteps, "ms_per_step": float(wt.item()) / wsteps,
		        "value": world * P * deg * wsteps / (float(wt.item()) * 1e-3), "unit": UNIT}  # fmt: skip

	# ---- end-to-end through the public API with host buffers ----------------------------------------------------
	e2e = None
	if not args.no_e2e:
		Ah = pinned_csr(A)
		h2d = Ah.data.nbytes + Ah.indices.nbytes + Ah.indptr.nbytes
		from primate_b200 import lanczos as _lz

		def e2e_step(s: int):
			_lz._OP_CACHE.clear()  # no operator caching across steps: the upload is part of every step
			Me = MatrixFunction(Ah, "log", deg=deg, orth=0)
			rng = np.random.default_rng(0)
			rng.bit_generator.advance(s * P * n)  # step s starts at probe s*P of the one stream
			return hutch(Me, converge="count", count=P, seed=rng)  # the P probes are sharded over the ranks by `comm="auto"`

		for s in range(min(args.warmup, 1)):
			e2e_step(s)
		sync_all()
		f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
		w0 = time.perf_counter()
		f0.record()
		est = 0.0
		for s in range(args.steps):
			est = e2e_step(args.warmup + s)
		f1.record()
		sync_all()
		wall = time.perf_counter() - w0
		ems = torch.tensor([max(f0.elapsed_time(f1), 0.0)], dtype=torch.float64, device="cuda")
		if world > 1:
			dist.all_reduce(ems, op=dist.ReduceOp.MAX)
		e2e = {"value": total_probe_steps / (float(ems.item()) * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
		       "d2h_bytes_per_step": int(8 * P if world == 1 else 24), "ms_per_step": float(ems.item()) / args.steps,
		       "wall_ms_per_step": 1e3 * wall / args.steps, "estimate": float(est),
		       "api": "hutch(MatrixFunction(A_host_csr, 'log', deg, orth=0), converge='count', count=P, seed=rng)",
		       "ingestion": "every step, every rank: new MatrixFunction from the pinned host CSR -> row permutation applied on the host "
		                    "(the locality ORDER itself is remembered per sparsity structure after the warm-up step) -> H2D of the CSR -> device operator"}  # fmt: skip

	if rank != 0:
		if world > 1:
			dist.destroy_process_group()
		return
	run_ours_report(args, world, ms, value, launches, clk, prof, e2e, rec, n, nnz, hi - lo, weak)
	if world > 1:
		dist.destroy_process_group()


def measured_peak_hbm() -> tuple:
	path = os.path.join(ROOT, "MEASURED_PEAKS.json")
	try:
		with open(path) as f:
			return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
	except (OSError, KeyError, ValueError):
		return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(key: str):
	"""DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture (profiles/), or None."""
	try:
		with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
			return json.load(f).get(key)
	except (OSError, ValueError):
		return None


def run_ours_report(args, world, ms, value, launches, clk, prof, e2e, rec, n, nnz, my_probes, weak) -> None:
	from primate_b200 import engine

	P, deg = args.probes, args.deg
	blocks = engine.block_sizes(my_probes, args.block)  # rank 0's share of the job
	csr_bytes = nnz * 12 + (n + 1) * 4
	# step-A algorithmic bytes per launch: CSR once + read u_c + read u_p (absent at j = 0) + write w  
	stepA_bytes = sum(deg * csr_bytes + (3 * deg - 1) * n * k * 8 for k in blocks) / (deg * len(blocks))
	axpy_bytes = sum(3 * n * k * 8 for k in blocks) / len(blocks)
	peak, peak_src = measured_peak_hbm()
	tA, cA, seenA = prof["stepA"]
	tB, cB, _ = prof["axpy"]
	avgA = tA / max(cA, 1)  # ms
	avgB = tB / max(cB, 1)
	# one kernel per Lanczos step (k_stepA_tma<STEP_FUSED>) for every block of >= 8 probes: (next to) no separate axpy launches
	# were seen (a rank's last block of < 8 probes still takes the generic two-kernel path: 1 block of 20 at N = 8)
	fused = cA > 0 and cB * 4 < cA
	if fused:
		# the fused step: CSR once + read the pending w and its vector u + write the formed vector and the new w; the first
		# step of a block reads the start block (twice, the second time from L2) and writes only w
		stepA_bytes = sum(deg * csr_bytes + ((4 * deg - 2) if k >= 8 else (3 * deg - 1)) * n * k * 8 for k in blocks) / (deg * len(blocks))
		kernel = ("k_stepA_tma<STEP_FUSED>: the WHOLE Lanczos step in one kernel (CSR SpMM on the vector formed on the fly from the "
		          "previous step's w and u, beta*q_p subtraction, alpha dot, norm sums; TMA-staged tiles) -- 4 panel passes per step "
		          "instead of the 6 of SURVEY 8d's two-kernel step")  # fmt: skip
		tkey = "k_stepA_fused_dram_bytes_per_launch"
	else:
		kernel = "k_stepA_tma (fused CSR SpMM + beta*q_p subtraction + alpha dot; TMA-staged tiles)"
		tkey = "k_stepA_dram_bytes_per_launch"
	achieved = stepA_bytes / (avgA * 1e-3) / 1e9 if cA else None
	roofline = {
		"bound": "hbm", "kernel": kernel,
		"achieved": achieved, "peak": peak, "unit": "GB/s", "frac": (achieved / peak) if achieved else None,
		"traffic": ncu_traffic(tkey), "peak_source": peak_src,
		"algorithmic_bytes_per_launch": stepA_bytes, "avg_launch_ms": avgA, "launches_timed": cA, "launches_seen": seenA,
		"share_of_step": (tA * seenA / max(cA, 1)) / (ms if ms > 0 else 1.0),
	}  # fmt: skip
	if fused:
		# the same launches against SURVEY 8(d)'s own per-unit figure (CSR/k + 6 n s per probe-step x the probes of a launch):
		# above 1 means the kernel beats the traffic model of the two-kernel step, not the hardware
		b8d = sum(deg * (csr_bytes + 6 * n * k * 8) for k in blocks) / (deg * len(blocks))
		roofline["survey_8d_accounting"] = {"algorithmic_bytes_per_launch": b8d, "achieved": b8d / (avgA * 1e-3) / 1e9 if cA else None,
		                                    "frac": (b8d / (avgA * 1e-3) / 1e9 / peak) if cA else None}  # fmt: skip
	else:
		roofline["axpy_kernel"] = {"avg_launch_ms": avgB, "algorithmic_bytes_per_launch": axpy_bytes,
		                           "achieved": (axpy_bytes / (avgB * 1e-3) / 1e9) if cB else None,
		                           "frac": (axpy_bytes / (avgB * 1e-3) / 1e9 / peak) if cB else None}  # fmt: skip
	# whole fused Lanczos step against SURVEY 8(d)'s B_step = CSR/k + 6 n s per probe-step
	bstep = sum(k * deg * (csr_bytes / k + 6 * n * 8) for k in blocks) / (my_probes * deg)
	roofline["whole_step"] = {"algorithmic_bytes_per_probe_step": bstep, "achieved": bstep * (value / world) / 1e9,
	                          "frac": bstep * (value / world) / 1e9 / peak}  # fmt: skip

	cpu = None
	if world == 1 and not args.no_cpu_baseline:
		cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "1", "--warmup", "1", "--grid", str(args.grid),
		       "--deg", str(deg), "--cpu-probes", str(args.cpu_probes or 3), "--cpu-workers", str(args.cpu_workers)]  # fmt: skip
		env = {k: v for k, v in os.environ.items() if k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE")}
		try:
			out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env)
			ref = json.loads([ln for ln in out.stdout.splitlines() if ln.startswith("{")][-1])
			cpu = ref["cpu_baseline"]
		except Exception as exc:  # noqa: BLE001  (the baseline is a reported extra; never fail the bench line over it)
			cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": f"failed: {exc!r}"}

	configs = None
	if world == 1 and not args.no_configs:
		configs = run_side_configs(args, peak)

	line = {
		"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
		"ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
		"data": "synthetic", "config": config_of(args),
		"details": {"block": args.block, "parallelism": f"probes of one {P}-probe job sharded x{world} (strong scaling)",
		            "probes_on_rank0": my_probes,
		            "l2_policy": "inputs larger than L2 (4 panels of n x block fp64 = 2 GB at block 64, + 64 MB CSR, vs 126 MB L2); no flush needed",
		            "collective": "one all-gather of (count, mean, M2) per rank per step" if world > 1 else "none"},
		"e2e": e2e, "gpu_launches": int(launches), "clocks": clk, "roofline": roofline, "cpu_baseline": cpu,
		"check": {"count": rec[0], "mean_logdet": rec[1], "std": float(np.sqrt(rec[2] / max(rec[0] - 1, 1)))},
	}  # fmt: skip
	if weak is not None:
		line["weak"] = weak
	if configs is not None:
		line["configs"] = configs
	print(json.dumps(line), flush=True)


def run_side_configs(args, peak_hbm: float) -> dict:
	"""BASELINE.json's other configs on this GPU, each at a stated size that finishes in well under a minute, through the
	generators and checks of scripts/run_configs.py: throughput in probe-steps/s and the fraction of the roofline that
	bounds the config (HBM bytes per probe-step `CSR/k + 6 n s` for the sparse ones; for the dense C4 the
	fp64 matrix product, against cuBLAS DGEMM timed here on the same shape as the yardstick)."""
	import types

	import torch

	from scripts import run_configs as rc

	out = {}
	a = types.SimpleNamespace(c3_scale=args.c3_scale, c3_probes=64, c3_block=64, c3_reorder="balance", c4_n=32768, c4_block=64, c5_grid=256, c5_small=48)

	def guarded(name, fn):
		t0 = time.perf_counter()
		try:
			res = fn()
		except Exception as exc:  # noqa: BLE001  (a side measurement never fails the bench line)
			res = {"error": repr(exc)}
		res["wall_s"] = time.perf_counter() - t0
		out[name] = res
		torch.cuda.empty_cache()

	def c1():
		r = rc.c1(a)
		return {"workload": "C1: hutch logdet, dense SPD 1000x1000, deg=20, 200 probes, fp64 (full size; launch-bound at this n)",
		        "probe_steps_per_s": r["probe_steps_per_s"], "ok": r["ok"],
		        "roofline": {"bound": "launch latency", "achieved": None, "peak": None, "frac": None}}  # fmt: skip

	def c3():
		r = rc.c3(a)
		n, nnz = r["n"], r["nnz"]
		bstep = (nnz * 8 + (n + 1) * 4) / r.get("block", 32) + 6 * n * 4
		ach = bstep * r["probe_steps_per_s"] / 1e9
		return {"workload": f"C3: tr exp(-tL), R-MAT scale {a.c3_scale} (n={n}, nnz={nnz}, max degree {r['max_degree']}), fp32, deg=30, {a.c3_probes} probes in blocks of {a.c3_block} (BASELINE size: scale 24)",
		        "probe_steps_per_s": r["probe_steps_per_s"], "ok": r["ok"],
		        "roofline": {"bound": "hbm", "achieved": ach, "peak": peak_hbm, "unit": "GB/s", "frac": ach / peak_hbm,
		                     "algorithmic_bytes_per_probe_step": bstep}}  # fmt: skip

	def c4():
		r = rc.c4(a)
		n = r["n"]
		flops = 2.0 * n * n  # per probe-step: the dense product dominates (re-orthogonalisation: 4 n j flops at step j)
		A = torch.randn((n, n), device="cuda", dtype=torch.float64)
		X = torch.randn((n, a.c4_block), device="cuda", dtype=torch.float64)
		for _ in range(2):
			torch.matmul(A, X)
		e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
		e0.record()
		for _ in range(5):
			torch.matmul(A, X)
		e1.record()
		torch.cuda.synchronize()
		dgemm_tf = 5 * 2.0 * n * n * a.c4_block / (e0.elapsed_time(e1) * 1e-3) / 1e12
		del A, X
		ach = flops * r["probe_steps_per_s"] / 1e12
		return {"workload": f"C4: numerical rank, dense Gram {n}x{n} (rank {r['rank']}), smoothstep, deg=100, orth=100, 64 probes in blocks of {a.c4_block}, fp64 (full size)",
		        "probe_steps_per_s": r["probe_steps_per_s"], "ok": r["ok"], "rank_estimate": r["mean"],
		        "roofline": {"bound": f"fp64 tensor (yardstick: cuBLAS DGEMM n x n x {a.c4_block} timed in this run)", "achieved": ach, "peak": dgemm_tf,
		                     "unit": "TFLOP/s", "frac": ach / dgemm_tf}}  # fmt: skip

	def c5():
		r = rc.c5(a)
		n, nnz = r["n"], r["nnz"]
		bstep = (nnz * 12 + (n + 1) * 4) / 16 + 6 * n * 8
		ach = bstep * r["probe_steps_per_s"] / 1e9
		return {"workload": f"C5: f=inv, 3-D 7-point Laplacian 256^3 (n={n}), deg=80, 64 probes in blocks of 16, fp64 (full size); block drivers on 48^3",
		        "probe_steps_per_s": r["probe_steps_per_s"], "ok": r["ok"], "funm_block_ms": r.get("funm_block_ms"),
		        "roofline": {"bound": "hbm", "achieved": ach, "peak": peak_hbm, "unit": "GB/s", "frac": ach / peak_hbm,
		                     "algorithmic_bytes_per_probe_step": bstep}}  # fmt: skip

	for name, fn in (("C1", c1), ("C3", c3), ("C4", c4), ("C5", c5)):
		guarded(name, fn)
	return out


def main():
	args = parse_args()
	if args.impl == "reference":
		run_reference(args)
	else:
		run_ours(args)


if __name__ == "__main__":
	main()
