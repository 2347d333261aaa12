"""Block drivers on several GPUs: `hutchpp`, `xdiag` (panels resident on the GPUs), `M @ block` (host panels) and `M.matmat_device` on a MatrixFunction built with
comm="auto" (columns of every block product split over the ranks, panels all-gathered over NCCL) against the same calls
on one GPU.  Launch:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
      scripts/multi_gpu_block_drivers.py [--grid 96] [--out gpurun_out/block_drivers.json]

Rank 0 prints one JSON line; exit code 1 if a sharded result differs from the single-GPU one by more than 1e-10."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def laplacian_3d(g: int):
	import scipy.sparse as sp

	T = sp.diags([-1.0, 2.0, -1.0], [-1, 0, 1], shape=(g, g))
	I = sp.identity(g)
	return (sp.kron(sp.kron(T, I), I) + sp.kron(sp.kron(I, T), I) + sp.kron(sp.kron(I, I), T)).tocsr()


def main():
	ap = argparse.ArgumentParser()
	ap.add_argument("--grid", type=int, default=96)
	ap.add_argument("--deg", type=int, default=80)
	ap.add_argument("--m", type=int, default=64)
	ap.add_argument("--only", default="matmat,matmat_device,xdiag,hutchpp")
	ap.add_argument("--out", default="gpurun_out/block_drivers.json")
	args = ap.parse_args()

	import torch
	import torch.distributed as dist

	local = int(os.environ.get("LOCAL_RANK", "0"))
	torch.cuda.set_device(local)
	os.environ.setdefault("NCCL_DEBUG", "WARN")
	dist.init_process_group("nccl", device_id=torch.device("cuda", local))
	rank, world = dist.get_rank(), dist.get_world_size()

	from primate_b200.diagonal import xdiag
	from primate_b200.operators import MatrixFunction
	from primate_b200.trace import hutchpp

	A = laplacian_3d(args.grid)
	n = A.shape[0]
	solo = MatrixFunction(A, "inv", deg=args.deg, orth=0)
	shard = MatrixFunction(A, "inv", deg=args.deg, orth=0, comm="auto")
	assert shard._comm.size == world

	def timed(fn):
		torch.cuda.synchronize()
		dist.barrier()
		t0 = time.perf_counter()
		y = fn()
		torch.cuda.synchronize()
		dist.barrier()
		return y, time.perf_counter() - t0

	X = np.random.default_rng(5).standard_normal((n, args.m))
	rec = {"n": n, "nnz": int(A.nnz), "deg": args.deg, "m": args.m, "world": world}
	names = args.only.split(",")
	Xd = torch.from_numpy(X).cuda()
	for name, call in (
		("matmat", lambda M: M @ X),
		("matmat_device", lambda M: M.matmat_device(Xd)),
		("xdiag", lambda M: xdiag(M, m=args.m, pdf="rademacher", seed=2)),
		("hutchpp", lambda M: np.array([hutchpp(M, m=96, seed=1)])),
	):
		if name not in names:
			continue
		call(shard)  # warm-up (workspaces, NCCL channels)
		y1, t1 = timed(lambda: call(solo))
		yN, tN = timed(lambda: call(shard))
		y1, yN = [y.cpu().numpy() if hasattr(y, "cpu") else y for y in (y1, yN)]
		err = float(np.max(np.abs(np.asarray(yN) - np.asarray(y1))) / np.max(np.abs(np.asarray(y1))))
		# the same bytes on every rank: compare a checksum
		chk = torch.tensor([float(np.sum(np.asarray(yN, dtype=np.float64)))], dtype=torch.float64, device="cuda")
		all_chk = [torch.zeros_like(chk) for _ in range(world)]
		dist.all_gather(all_chk, chk)
		same = all(float(c.item()) == float(all_chk[0].item()) for c in all_chk)
		rec[name] = {"rel_err_vs_one_gpu": err, "identical_on_all_ranks": same, "seconds_one_gpu": t1, "seconds_sharded": tN,
		             "speedup": t1 / tN}  # fmt: skip
	rec["ok"] = all(rec[k]["rel_err_vs_one_gpu"] < 1e-10 and rec[k]["identical_on_all_ranks"] for k in names)
	if rank == 0:
		os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
		with open(args.out, "w") as f:
			f.write(json.dumps(rec) + "\n")
		print(json.dumps(rec), flush=True)
	dist.barrier()
	dist.destroy_process_group()
	return 0 if rec["ok"] else 1


if __name__ == "__main__":
	sys.exit(main())
