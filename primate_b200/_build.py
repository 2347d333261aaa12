"""In-tree nvcc build of libslq.so (sm_100a only).  `python -m primate_b200._build [--force]`."""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
INCLUDE = os.path.join(ROOT, "include")
OBJ_DIR = os.path.join(PKG, "csrc", "build")
LIB = os.path.join(PKG, "libslq.so")
SOURCES = ["slq_api.cu", "slq_lanczos.cu", "slq_probes.cu", "slq_quadrature.cu", "slq_reorder.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
	"-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
	"-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--expt-relaxed-constexpr",
	"-I" + INCLUDE, "-I" + CSRC, "-DSLQ_BUILD",
]  # fmt: skip
FLAGS += os.environ.get("SLQ_EXTRA_NVCC_FLAGS", "").split()  # tuning experiments only (e.g. -DSLQ_STEPA_MINB=2)
LIB = os.environ.get("SLQ_LIB_OUT", LIB)


def _stale(target: str, deps: list) -> bool:
	if not os.path.exists(target):
		return True
	t = os.path.getmtime(target)
	return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def get_include() -> str:
	"""Directory of `slq.h` (the analogue of primate.get_include(), src/primate/__init__.py:17-41)."""
	return INCLUDE


def build(force: bool = False, verbose: bool = True) -> str:
	headers = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith(".cuh")] + [os.path.join(INCLUDE, "slq.h")]
	os.makedirs(OBJ_DIR, exist_ok=True)
	jobs = []
	for src in SOURCES:
		s, o = os.path.join(CSRC, src), os.path.join(OBJ_DIR, src.replace(".cu", ".o"))
		if force or _stale(o, [s] + headers):
			jobs.append([NVCC, *FLAGS, "-Xptxas", "-v", "-c", s, "-o", o])
	if not jobs and not _stale(LIB, [os.path.join(OBJ_DIR, s.replace(".cu", ".o")) for s in SOURCES]):
		return LIB
	if not os.path.exists(NVCC):
		raise RuntimeError(f"nvcc not found at {NVCC}; libslq.so cannot be built and there is no CPU fallback")

	def run(cmd):
		if verbose:
			print("[build] " + " ".join(cmd), flush=True)
		res = subprocess.run(cmd, capture_output=True, text=True)
		log = os.path.join(OBJ_DIR, os.path.basename(cmd[-1]) + ".log")
		with open(log, "w") as f:
			f.write(res.stdout + res.stderr)
		if res.returncode != 0:
			raise RuntimeError(f"nvcc failed:\n{res.stdout}\n{res.stderr}")

	with ThreadPoolExecutor(max_workers=4) as ex:
		list(ex.map(run, jobs))
	objs = [os.path.join(OBJ_DIR, s.replace(".cu", ".o")) for s in SOURCES]
	run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC", *objs, "-o", LIB])
	return LIB


if __name__ == "__main__":
	print(build(force="--force" in sys.argv))
