"""What the HBM of this box delivers per access mix, as yardsticks for the rooflines in DESIGN.md: MEASURED_PEAKS.json's
`hbm_gbs` is a COPY (half reads, half writes); the dense product reads 8.6 GB and writes 8 MB, step A reads 2/3 and
writes 1/3.  Library kernels only (torch reductions / fills / copies over 8 GiB, CUDA events, best of `--reps`) --
nothing here is on the product path.
  python scripts/bench_hbm_modes.py [--gib 8]"""
import argparse
import json


def main():
	ap = argparse.ArgumentParser()
	ap.add_argument("--gib", type=float, default=8.0)
	ap.add_argument("--reps", type=int, default=10)
	args = ap.parse_args()
	import torch

	nel = int(args.gib * (1 << 30)) // 8
	a = torch.ones(nel, dtype=torch.float64, device="cuda")
	b = torch.empty_like(a)
	a32 = a.view(torch.float32)

	def best(fn):
		for _ in range(2):
			fn()
		torch.cuda.synchronize()
		t = []
		for _ in range(args.reps):
			e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
			e0.record()
			fn()
			e1.record()
			torch.cuda.synchronize()
			t.append(e0.elapsed_time(e1))
		return min(t)

	nbytes = nel * 8
	out = {"gib": args.gib}
	for name, fn, moved in (
		("read_sum_f64", lambda: a.sum(), nbytes),
		("read_sum_f32", lambda: a32.sum(), nbytes),
		("read_max_f64", lambda: a.max(), nbytes),
		("read_dot_f64", lambda: torch.dot(a[: nel // 2], a[nel // 2 :]), nbytes),
		("write_fill_f64", lambda: b.fill_(1.0), nbytes),
		("write_zero", lambda: b.zero_(), nbytes),
		("copy", lambda: b.copy_(a), 2 * nbytes),
		("read2_write1_add", lambda: torch.add(a[: nel // 2], b[: nel // 2], out=b[nel // 2 :]), 3 * (nbytes // 2)),
	):
		ms = best(fn)
		out[name] = {"ms": round(ms, 4), "GBps": round(moved / ms / 1e6, 1)}
	print(json.dumps(out))


if __name__ == "__main__":
	main()
