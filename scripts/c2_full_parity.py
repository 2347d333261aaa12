"""Full-size parity of the headline job (config C2: 2-D Laplacian g=1000, f=log, deg=50, Rademacher probes, seed 0):
the log-det estimate and every per-probe quadrature sample of `hutch` on the GPU against the reference's compiled
extension (oracle/_ref) run on the same probe indices, sharded over the host cores (last row).

  python scripts/c2_full_parity.py [--probes 10000] [--out gpurun_out/c2_full_parity.json]

Test infrastructure (it executes oracle/): never imported by the package.  About 15 minutes on 16 host cores at the full
10 000 probes."""
import argparse
import json
import multiprocessing as mp
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (worker functions of the reference arm)


def main():
	ap = argparse.ArgumentParser()
	ap.add_argument("--probes", type=int, default=10000)
	ap.add_argument("--grid", type=int, default=1000)
	ap.add_argument("--deg", type=int, default=50)
	ap.add_argument("--task", type=int, default=4, help="probes per oracle task")
	ap.add_argument("--out", default="gpurun_out/c2_full_parity.json")
	args = ap.parse_args()

	import torch

	from primate_b200.operators import MatrixFunction
	from primate_b200.trace import hutch

	A = bench.laplacian_2d(args.grid)
	M = MatrixFunction(A, "log", deg=args.deg, orth=0)
	torch.cuda.synchronize()
	t0 = time.perf_counter()
	est, info = hutch(M, converge="count", count=args.probes, seed=0, full=True, record=True)
	torch.cuda.synchronize()
	gpu_s = time.perf_counter() - t0
	ours_all = np.concatenate([np.ravel(np.asarray(v, dtype=np.float64)) for v in info.estimator.values])
	ours = ours_all[: args.probes]  # the batched driver rounds the count up to whole batches of 32, as the reference does
	est_fixed = hutch(M, converge="count", count=args.probes, seed=0)  # the plain call: exactly `probes` samples

	workers = bench.host_cores()
	tasks = [(f, min(args.task, args.probes - f)) for f in range(0, args.probes, args.task)]
	t0 = time.perf_counter()
	with mp.get_context("spawn").Pool(workers, initializer=bench._ref_worker_init, initargs=(args.grid, args.deg)) as pool:
		out = pool.map(bench._ref_worker_run, tasks, chunksize=1)
	cpu_s = time.perf_counter() - t0
	ref = np.concatenate([o[0] for o in out])[: args.probes]

	rel = np.abs(ours / ref - 1)
	rec = {
		"config": "C2", "n": int(A.shape[0]), "probes": args.probes, "deg": args.deg, "seed": 0,
		"estimate_gpu": float(est_fixed), "estimate_ref": float(ref.mean()), "estimate_rel_err": float(abs(est_fixed / ref.mean() - 1)),
		"estimate_gpu_batched": float(est), "samples_batched": int(len(ours_all)),
		"batched_estimate_vs_own_samples_rel_err": float(abs(est / ours_all.mean() - 1)),
		"sample_rel_err_max": float(rel.max()), "sample_rel_err_median": float(np.median(rel)),
		"samples_compared": int(len(ref)), "gpu_seconds": gpu_s, "ref_seconds": cpu_s, "ref_workers": workers,
		"ref_probe_steps_per_s": args.probes * args.deg / cpu_s, "gpu_probe_steps_per_s_incl_setup": args.probes * args.deg / gpu_s,
		"ok": bool(abs(est_fixed / ref.mean() - 1) < 1e-8 and rel.max() < 1e-8 and abs(est / ours_all.mean() - 1) < 1e-12),
	}  # fmt: skip
	os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
	with open(args.out, "w") as f:
		f.write(json.dumps(rec) + "\n")
	print(json.dumps(rec))
	return 0 if rec["ok"] else 1


if __name__ == "__main__":
	sys.exit(main())
