"""Summarise a gpurun_out/<tag>/ capture (scripts/gpu_round.sh) into profiles/<tag>_*: the bench lines, the per-kernel
share of the ncu launch list and the key raw metrics of the --set full capture.  Usage: python scripts/summarize_ncu.py <tag>"""
import collections
import csv
import io
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
src = os.path.join(ROOT, "gpurun_out", tag)
dst = os.path.join(ROOT, "profiles")
os.makedirs(dst, exist_ok=True)

for f in ("bench.json", "bench_ref.json", "gpu.csv", "pytest_gpu.log", "smoke.log"):
	if os.path.exists(os.path.join(src, f)):
		shutil.copy(os.path.join(src, f), os.path.join(dst, f"{tag}_{f}"))

# ---- launch list: per-kernel totals and shares -------------------------------------------------------------------
lp = os.path.join(src, "launches.csv")
if os.path.exists(lp):
	lines = [ln for ln in open(lp) if ln.startswith('"')]
	rows = list(csv.DictReader(io.StringIO("".join(lines))))
	tot = collections.defaultdict(lambda: [0, 0.0])
	for r in rows:
		if r.get("Metric Name") != "gpu__time_duration.sum":
			continue
		name = r["Kernel Name"].split("(")[0]
		v = float(r["Metric Value"].replace(",", ""))
		unit = r.get("Metric Unit", "ns")
		v_us = v / 1e3 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1e3)
		tot[name][0] += 1
		tot[name][1] += v_us
	total = sum(t for _, t in tot.values())
	with open(os.path.join(dst, f"{tag}_launches_summary.md"), "w") as f:
		f.write(f"# ncu launch list ({tag}): `ncu --metrics gpu__time_duration.sum --clock-control none` over "
		        f"{sum(c for c, _ in tot.values())} launches inside the timed step (cold-cache, serialised: compare SHARES)\n\n")
		f.write("| kernel | launches | total us | avg us | share |\n|---|---|---|---|---|\n")
		for name, (c, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
			f.write(f"| `{name}` | {c} | {t:.1f} | {t / c:.2f} | {100 * t / total:.1f}% |\n")
	shutil.copy(lp, os.path.join(dst, f"{tag}_launches.csv"))

# ---- full capture: key raw metrics per profiled launch ---------------------------------------------------------------
rep = os.path.join(src, "prof_step.ncu-rep")
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"]  # fmt: skip
if os.path.exists(rep):
	out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
	rows = list(csv.reader(io.StringIO(out)))
	hdr, units = rows[0], rows[1]
	traffic = {}
	with open(os.path.join(dst, f"{tag}_ncu_full_summary.md"), "w") as f:
		f.write(f"# ncu --set full ({tag}), per profiled launch (`ncu -i prof_step.ncu-rep --page raw --csv`)\n")
		for r in rows[2:]:
			name = r[hdr.index("Kernel Name")].split("(")[0]
			f.write(f"\n## `{name}`\n\n| metric | value | unit |\n|---|---|---|\n")
			vals = {}
			for k in KEYS:
				if k in hdr:
					i = hdr.index(k)
					f.write(f"| {k} | {r[i]} | {units[i]} |\n")
					vals[k] = (r[i], units[i])
			try:
				def tobytes(k):
					v, u = vals[k]
					v = float(v.replace(",", ""))
					return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
				tb = tobytes("dram__bytes_read.sum") + tobytes("dram__bytes_write.sum")
				f.write(f"| **dram traffic (read+write)** | {tb / 1e6:.1f} | MB |\n")
				fused = "stepA" in name and ", 4, " in name  # k_stepA_tma<T, K, STEP_FUSED = 4, LONGROWS>
				key = ("k_stepA_fused" if fused else "k_stepA") if "stepA" in name else ("k_axpy" if "axpy" in name else name)
				traffic.setdefault(f"{key}_dram_bytes_per_launch", tb)
			except (KeyError, ValueError):
				pass
	tp = os.path.join(dst, "traffic.json")
	try:  # keep the entries of kernels this capture did not profile (e.g. the two-kernel step of an earlier round)
		with open(tp) as f:
			prev = json.load(f)
	except (OSError, ValueError):
		prev = {}
	prev.update(traffic)
	prev["source"] = (prev.get("source", "") + f"; profiles/{tag}_ncu_full_summary.md").lstrip("; ")
	with open(tp, "w") as f:
		json.dump(prev, f, indent=1)
print("profiles/ updated from", src)
