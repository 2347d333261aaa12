"""Build recipe for the two CPU checkers under ``oracle/`` (TEST INFRASTRUCTURE ONLY).

1. ``oracle/_ref/_lanczos*.so`` -- the reference's own pybind11 extension, compiled *in place* from
   ``/root/reference/src/primate/_lanczos.cpp`` (+ its headers and the vendored Eigen under
   ``/root/reference/extern/eigen``).  No reference source is copied into this repository: only the
   resulting binary lands in ``oracle/_ref/`` (git-ignored, NOT gpurun-ignored, so it travels to the
   GPU box).  The reference's meson build is not usable (``meson.build:72`` references a missing
   ``openmp`` subdir and undeclared options) so the single translation unit is compiled directly.
2. ``oracle/_build/liboracle_slq.so`` -- our plain-C restatement (``oracle/slq_oracle.c``).

Only ``tests/``, ``__graft_entry__`` and ``bench.py``'s cpu_baseline / ``--impl reference`` legs may
use either artefact.  The product package (``primate_b200``) never imports anything from here.
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("PRIMATE_REFERENCE_ROOT", "/root/reference")
REF_DIR = os.path.join(HERE, "_ref")
BUILD_DIR = os.path.join(HERE, "_build")
EXT_SUFFIX = sysconfig.get_config_var("EXT_SUFFIX")
REF_SO = os.path.join(REF_DIR, "_lanczos" + EXT_SUFFIX)
ORACLE_SO = os.path.join(BUILD_DIR, "liboracle_slq.so")


def _newer(target: str, *sources: str) -> bool:
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources if os.path.exists(s))


def build_reference_extension(force: bool = False, verbose: bool = True) -> str | None:
    """Compiles the reference's `_lanczos` extension; returns the path or None if the sources are absent."""
    src = os.path.join(REF_ROOT, "src", "primate", "_lanczos.cpp")
    if not os.path.exists(src):
        return REF_SO if os.path.exists(REF_SO) else None  # GPU box: prebuilt binary only
    if not force and _newer(REF_SO, src):
        return REF_SO
    import pybind11

    os.makedirs(REF_DIR, exist_ok=True)
    cmd = [
        "g++", "-O3", "-DNDEBUG", "-std=c++20", "-shared", "-fPIC", "-fvisibility=hidden", "-w",
        "-I" + os.path.join(REF_ROOT, "src", "primate", "include"),
        "-I" + os.path.join(REF_ROOT, "extern", "eigen"),
        "-I" + pybind11.get_include(),
        "-I" + sysconfig.get_paths()["include"],
        src, "-o", REF_SO,
    ]  # fmt: skip
    if verbose:
        print("[oracle] " + " ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    return REF_SO


def build_c_oracle(force: bool = False, verbose: bool = True) -> str:
    """Compiles our C restatement of the path (oracle/slq_oracle.c)."""
    src = os.path.join(HERE, "slq_oracle.c")
    if not force and _newer(ORACLE_SO, src):
        return ORACLE_SO
    os.makedirs(BUILD_DIR, exist_ok=True)
    # -ffp-contract=off: no FMA contraction, so the C port rounds like the reference's Eigen code paths
    cmd = ["gcc", "-O2", "-ffp-contract=off", "-std=c11", "-shared", "-fPIC", src, "-o", ORACLE_SO, "-lm"]
    if verbose:
        print("[oracle] " + " ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    return ORACLE_SO


def build_all(force: bool = False, verbose: bool = True) -> dict:
    return {"reference_ext": build_reference_extension(force, verbose), "c_oracle": build_c_oracle(force, verbose)}


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv))
