"""In-tree nvcc build of libslq.so (sm_100a only).  `python -m primate_b200._build [--force]`.

Staleness is decided by CONTENT, not by mtimes: every object file is keyed on the SHA-256 of its source, of all the
library's headers and of the compiler flags, and the shared library on its objects' keys (`csrc/build/manifest.json`
and `libslq.so.manifest` travel with the binaries).  A prebuilt library copied to another machine (the GPU box gets a
snapshot of this tree with fresh mtimes) is therefore loaded as is, and a library that no longer matches the sources
is never loaded silently: it is rebuilt, or -- without nvcc -- the import fails.
"""
from __future__ import annotations

import hashlib
import json
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
SOURCES = [
	"slq_api.cu", "slq_lanczos.cu", "slq_lanczos_f64.cu", "slq_lanczos_f32.cu", "slq_probes.cu", "slq_normals.cu", "slq_panels.cu", "slq_quadrature.cu",
	"slq_reorder.cu",
]  # fmt: skip
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
	"-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
	"-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--expt-relaxed-constexpr",
	"-I" + INCLUDE, "-I" + CSRC, "-DSLQ_BUILD",
]  # fmt: skip
FLAGS += os.environ.get("SLQ_EXTRA_NVCC_FLAGS", "").split()  # tuning experiments only (e.g. -DSLQ_STEPA_MINB=2)
LIB = os.environ.get("SLQ_LIB_OUT", LIB)
MANIFEST = os.path.join(OBJ_DIR, "manifest.json")


def get_include() -> str:
	"""Directory of `slq.h` (the analogue of primate.get_include(), src/primate/__init__.py:17-41)."""
	return INCLUDE


def _sha(paths: list, extra: str = "") -> str:
	h = hashlib.sha256(extra.encode())
	for p in paths:
		h.update(os.path.basename(p).encode())
		with open(p, "rb") as f:
			h.update(f.read())
	return h.hexdigest()


def _headers() -> list:
	return [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cuh", ".h"))] + [os.path.join(INCLUDE, "slq.h")]


def source_keys() -> dict:
	"""{object file name: content key} for the current sources, headers and flags."""
	hdr = _sha(_headers(), " ".join(f for f in FLAGS if not f.startswith("-I")))
	return {src.replace(".cu", ".o"): _sha([os.path.join(CSRC, src)], hdr) for src in SOURCES}


def _read_json(path: str) -> dict:
	try:
		with open(path) as f:
			return json.load(f)
	except (OSError, ValueError):
		return {}


def is_current() -> bool:
	"""True when libslq.so exists and was built from exactly the sources in this tree."""
	return os.path.exists(LIB) and _read_json(LIB + ".manifest") == source_keys()


def build(force: bool = False, verbose: bool = True) -> str:
	keys = source_keys()
	if not force and os.path.exists(LIB) and _read_json(LIB + ".manifest") == keys:
		return LIB
	if not os.path.exists(NVCC):
		raise RuntimeError(f"libslq.so is missing or stale and nvcc was not found at {NVCC}; there is no CPU fallback")
	os.makedirs(OBJ_DIR, exist_ok=True)
	have = {} if force else _read_json(MANIFEST)
	jobs = []
	for src in SOURCES:
		o = src.replace(".cu", ".o")
		if have.get(o) != keys[o] or not os.path.exists(os.path.join(OBJ_DIR, o)):
			jobs.append([NVCC, *FLAGS, "-Xptxas", "-v", "-c", os.path.join(CSRC, src), "-o", os.path.join(OBJ_DIR, o)])

	def run(cmd):
		if verbose:
			print("[build] " + " ".join(cmd), flush=True)
		res = subprocess.run(cmd, capture_output=True, text=True)
		log = os.path.join(OBJ_DIR, os.path.basename(cmd[-1]) + ".log")
		with open(log, "w") as f:
			f.write(res.stdout + res.stderr)
		if res.returncode != 0:
			raise RuntimeError(f"nvcc failed:\n{res.stdout}\n{res.stderr}")

	with ThreadPoolExecutor(max_workers=int(os.environ.get("SLQ_BUILD_JOBS", "8"))) as ex:
		list(ex.map(run, jobs))
	for cmd in jobs:
		o = os.path.basename(cmd[-1])
		have[o] = keys[o]
	with open(MANIFEST, "w") as f:
		json.dump({o: have[o] for o in keys}, f, indent=1)
	objs = [os.path.join(OBJ_DIR, s.replace(".cu", ".o")) for s in SOURCES]
	run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC", *objs, "-o", LIB])
	with open(LIB + ".manifest", "w") as f:
		json.dump(keys, f, indent=1)
	return LIB


if __name__ == "__main__":
	print(build(force="--force" in sys.argv))
