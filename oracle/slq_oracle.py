"""CPU oracle for the stochastic-Lanczos-quadrature path of peekxc/primate -- TEST INFRASTRUCTURE ONLY.

This module is the *checker* for the CUDA library: only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import it.  ``primate_b200`` (the
product) never does, and raises if its CUDA extension is missing instead of falling back to anything
in here.

Three layers, each citing the reference file:line (relative to /root/reference) it follows:

* ``ref_*``  -- thin harness around the **reference's own compiled extension** (``oracle/_ref/_lanczos*.so``,
  built from /root/reference/src/primate/_lanczos.cpp by ``oracle/build_oracle.py``; binary only, no
  reference sources live in this repo).  The Python glue the reference wraps around it
  (``operators.py:126-151``, ``integrate.py:58-64``, ``trace.py:86-116``) is restated here in our own words.
* ``np_*``   -- NumPy restatement of the same algorithms (small cases / second opinion).
* ``c_*``    -- ctypes bindings of the plain-C restatement ``oracle/slq_oracle.c``.

Parity is PINNED: ``tests/test_oracle.py`` checks all three layers against each other and against the
golden vectors in ``tests/golden/`` which were produced by importing the unmodified reference package
in the build container (``tests/golden/make_golden.py``).

Third-party arithmetic the reference delegates to (not under /root/reference), used here verbatim as the
reference does: NumPy 2.3.5 ``Generator(PCG64)`` (random.py:23,33) and SciPy 1.18.1
``scipy.linalg.eigh_tridiagonal`` -> LAPACK ``dstemr`` (tridiag.py:3,10-11).
"""
from __future__ import annotations

import ctypes
import importlib.util
import os
from typing import Callable, Optional, Union

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_REF_MOD = None
_C_LIB = None

PCG_MULT = 0x2360ED051FC65DA44385DF649FCCF645
MASK128 = (1 << 128) - 1
MASK64 = (1 << 64) - 1


# --------------------------------------------------------------------------------------------------
# loaders
# --------------------------------------------------------------------------------------------------
def reference_ext():
    """The reference's pybind11 module `_lanczos` (src/primate/_lanczos.cpp:102-112), prebuilt in oracle/_ref."""
    global _REF_MOD
    if _REF_MOD is None:
        import glob

        cands = sorted(glob.glob(os.path.join(HERE, "_ref", "_lanczos*.so")))
        if not cands:
            raise FileNotFoundError("oracle/_ref/_lanczos*.so missing: run `python oracle/build_oracle.py`")
        spec = importlib.util.spec_from_file_location("_lanczos", cands[0])
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _REF_MOD = mod
    return _REF_MOD


def have_reference_ext() -> bool:
    try:
        reference_ext()
        return True
    except Exception:
        return False


class _OrcOp(ctypes.Structure):
    _fields_ = [
        ("kind", ctypes.c_int),
        ("n", ctypes.c_int64),
        ("indptr", ctypes.c_void_p),
        ("indices", ctypes.c_void_p),
        ("data", ctypes.c_void_p),
    ]


def c_lib():
    global _C_LIB
    if _C_LIB is None:
        path = os.path.join(HERE, "_build", "liboracle_slq.so")
        if not os.path.exists(path):
            raise FileNotFoundError("oracle/_build/liboracle_slq.so missing: run `python oracle/build_oracle.py`")
        lib = ctypes.CDLL(path)
        u64p = ctypes.POINTER(ctypes.c_uint64)
        lib.orc_pcg64_raw.argtypes = [u64p, u64p, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int64, ctypes.c_void_p]
        lib.orc_rademacher.argtypes = [u64p, u64p, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int64, ctypes.c_void_p]
        for suf, ft in (("f64", ctypes.c_double), ("f32", ctypes.c_float)):
            fn = getattr(lib, "orc_lanczos_" + suf)
            fn.argtypes = [ctypes.POINTER(_OrcOp), ctypes.c_void_p, ctypes.c_int, ft, ctypes.c_int,
                           ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]  # fmt: skip
            fn.restype = ctypes.c_int
        lib.orc_tqli.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
        lib.orc_fttr.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        lib.orc_slq_quad_f64.argtypes = [ctypes.POINTER(_OrcOp), ctypes.c_void_p, ctypes.c_int, ctypes.c_double,
                                         ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                         ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]  # fmt: skip
        lib.orc_slq_quad_f64.restype = ctypes.c_double
        _C_LIB = lib
    return _C_LIB


# --------------------------------------------------------------------------------------------------
# probes: src/primate/random.py:22-29,47-80 on NumPy's PCG64
# --------------------------------------------------------------------------------------------------
def pcg64_state(seed) -> tuple:
    """(state, inc) 128-bit ints of ``np.random.default_rng(seed)`` *before* any draw (trace.py:86)."""
    bg = seed.bit_generator if isinstance(seed, np.random.Generator) else np.random.PCG64(seed)
    st = bg.state
    assert st["bit_generator"] == "PCG64"
    return int(st["state"]["state"]), int(st["state"]["inc"])


def np_pcg64_raw(state: int, inc: int, offset: int, count: int) -> np.ndarray:
    """Pure-Python PCG64 XSL-RR 128/64: raw outputs #offset..#offset+count-1 (small counts only)."""
    # O(log offset) jump-ahead (Brown 1994), same recurrence PCG64.advance uses
    acc_m, acc_p, cur_m, cur_p, d = 1, 0, PCG_MULT, inc, offset
    while d > 0:
        if d & 1:
            acc_m = (acc_m * cur_m) & MASK128
            acc_p = (acc_p * cur_m + cur_p) & MASK128
        cur_p = ((cur_m + 1) * cur_p) & MASK128
        cur_m = (cur_m * cur_m) & MASK128
        d >>= 1
    s = (acc_m * state + acc_p) & MASK128
    out = np.empty(count, dtype=np.uint64)
    for i in range(count):
        s = (s * PCG_MULT + inc) & MASK128  # step ...
        hi, lo = s >> 64, s & MASK64
        x, rot = hi ^ lo, hi >> 58
        out[i] = ((x >> rot) | (x << ((64 - rot) & 63))) & MASK64  # ... then output
    return out


def c_rademacher(state: int, inc: int, offset: int, count: int) -> np.ndarray:
    st = (ctypes.c_uint64 * 2)(state >> 64, state & MASK64)
    ic = (ctypes.c_uint64 * 2)(inc >> 64, inc & MASK64)
    out = np.empty(count, dtype=np.float64)
    c_lib().orc_rademacher(st, ic, offset >> 64, offset & MASK64, count, out.ctypes.data)
    return out


def c_pcg64_raw(state: int, inc: int, offset: int, count: int) -> np.ndarray:
    st = (ctypes.c_uint64 * 2)(state >> 64, state & MASK64)
    ic = (ctypes.c_uint64 * 2)(inc >> 64, inc & MASK64)
    out = np.empty(count, dtype=np.uint64)
    c_lib().orc_pcg64_raw(st, ic, offset >> 64, offset & MASK64, count, out.ctypes.data)
    return out


def ref_isotropic(size: tuple, pdf: str = "rademacher", seed=None) -> np.ndarray:
    """random.py:47-80 restated on the same NumPy Generator calls the reference makes (F-ordered fill)."""
    rng = np.random.default_rng(seed)
    size = (size, 1) if isinstance(size, int) else size
    W = np.empty(shape=size, dtype=np.float64, order="F")
    if pdf in ("rademacher", "signs"):
        rng.random(out=W)  # random.py:23
        np.multiply(W, 2, out=W)
        np.floor(W, out=W)
        np.multiply(W, 2, out=W)
        np.subtract(W, 1, out=W)  # random.py:26-29
    elif pdf in ("normal", "gaussian"):
        rng.standard_normal(out=W, dtype=W.dtype)  # random.py:33
    elif pdf == "sphere":
        rng.standard_normal(out=W, dtype=W.dtype)  # random.py:37-41
        c = np.sqrt(np.sum(W**2, axis=0, keepdims=True))
        np.divide(W, c, out=W)
        np.multiply(W, np.sqrt(W.shape[0]), out=W)
    else:
        raise ValueError(pdf)
    return W


# --------------------------------------------------------------------------------------------------
# spectral functions: src/primate/special.py:78-107
# --------------------------------------------------------------------------------------------------
def spectral_fun(fun: Union[str, Callable, None], **kw) -> Callable:
    if fun is None or fun == "identity":
        return lambda x: x
    if callable(fun):
        return fun
    if fun == "abs":
        return np.abs
    if fun == "sqrt":
        return np.sqrt
    if fun == "log":
        return lambda x: np.log(np.maximum(x, np.finfo(np.float64).eps))  # special.py:90
    if fun == "inv":
        return np.reciprocal
    if fun == "exp":
        t = kw.get("t", 1.0)
        return lambda x: np.exp(t * x)  # special.py:62-66
    if fun == "smoothstep":
        a, b = kw.get("a", 0.0), kw.get("b", 1.0)
        d = (b - a) if a != b else 1.0

        def _ss(x):  # special.py:46-49
            y = np.clip((x - a) / d, 0.0, 1.0)
            return 3 * y**2 - 2 * y**3

        return _ss
    if fun == "numrank":
        c = kw.get("threshold", 0.000001)
        return lambda x: np.where(np.abs(x) < c, 0.0, 1.0)  # special.py:69-74,104-105
    raise ValueError(f"Unknown function: {fun}.")


# --------------------------------------------------------------------------------------------------
# Lanczos: include/lanczos.h:43-66, 92-149
# --------------------------------------------------------------------------------------------------
def _ncv_for(deg: int, orth: int, return_basis: bool = False) -> tuple:
    """lanczos.py:88-89 parameter policy -> (orth, ncv)."""
    orth = deg if orth < 0 or orth > deg else orth
    ncv = int(np.clip(orth, 2, deg)) if not return_basis else deg
    return orth, ncv


def ref_lanczos(A, v: np.ndarray, deg: int, rtol: float = 1e-8, orth: int = 0, ncv: Optional[int] = None,
                Q: Optional[np.ndarray] = None):
    """One call of the reference FFI (src/primate/_lanczos.cpp:86-100) on FRESH zero buffers
    (lanczos.py:101-103).  Returns (alpha[deg+1], beta[deg+1], Q[n, ncv]).

    ``Q`` may be a caller-owned (n, ncv) F-ordered buffer that is re-used as is, the way the reference's
    ``MatrixFunction`` re-uses ``_Q`` across probes (operators.py:138-148).  Only valid for ``orth == 0``, where the
    result does not depend on what the stale columns hold; for ``orth > 0`` it is zeroed first."""
    ext = reference_ext()
    n = A.shape[0]
    dt = np.dtype(v.dtype)
    if ncv is None:
        orth, ncv = _ncv_for(deg, orth)
    alpha = np.zeros(deg + 1, dtype=dt)
    beta = np.zeros(deg + 1, dtype=dt)
    if Q is None:
        Q = np.zeros((n, ncv), dtype=dt, order="F")
    else:
        assert Q.shape == (n, ncv) and Q.dtype == dt and Q.flags["F_CONTIGUOUS"]
        if orth > 0:
            Q.fill(0)
    ext.lanczos(A, np.array(v, dtype=dt, copy=True), deg, rtol, orth, alpha, beta, Q)
    return alpha, beta, Q


def np_orth_vector(v, U, start_idx, p, reverse=False):
    """lanczos.h:43-66 in NumPy (in place on v)."""
    n, m = U.shape
    tol = 2 * np.finfo(U.dtype).eps * np.sqrt(n)
    diff = -1 if reverse else 1
    i = start_idx % m
    for _ in range(p):
        u = U[:, i]
        u_norm = np.dot(u, u)
        s_proj = np.dot(v, u)
        if u_norm > tol and abs(s_proj) > tol:
            v -= (s_proj / u_norm) * u
        i = (i + diff) % m


def np_lanczos(A, v: np.ndarray, deg: int, rtol: float = 1e-8, orth: int = 0, ncv: Optional[int] = None):
    """lanczos.h:92-149 in NumPy.  Returns (alpha[deg+1], beta[deg+1], Q[n, ncv], steps)."""
    n = A.shape[0]
    dt = np.dtype(v.dtype)
    if ncv is None:
        orth, ncv = _ncv_for(deg, orth)
    residual_tol = dt.type(np.sqrt(n)) * dt.type(rtol)
    alpha, beta = np.zeros(deg + 1, dtype=dt), np.zeros(deg + 1, dtype=dt)
    Q = np.zeros((n, ncv), dtype=dt, order="F")
    pos = [ncv - 1, 0, 1]
    Q[:, pos[0]] = 0
    nrm = np.sqrt(np.dot(v, v))
    Q[:, 0] = v / nrm if nrm > 0 else v
    steps = 0
    for j in range(deg):
        p, c, nx = pos
        w = np.asarray(A @ Q[:, c], dtype=dt).ravel()
        w -= beta[j] * Q[:, p]
        alpha[j] = np.dot(Q[:, c], w)
        w -= alpha[j] * Q[:, c]
        if orth > 0:
            np_orth_vector(w, Q, c, orth, reverse=True)
        beta[j + 1] = np.sqrt(np.dot(w, w))
        steps = j + 1
        if beta[j + 1] < residual_tol or (j + 1) == deg:
            break
        Q[:, nx] = w / beta[j + 1]
        pos = [c, nx, (j + 2) % ncv]
    return alpha, beta, Q, steps


def _make_orc_op(A, dt):
    """CSC arrays (kept alive in the returned tuple) for the C port; dense -> col-major copy."""
    import scipy.sparse as sp

    if sp.issparse(A):
        C = sp.csc_matrix(A)  # what the pybind11 sparse caster does
        C.sort_indices()
        indptr = np.ascontiguousarray(C.indptr, dtype=np.int64)
        indices = np.ascontiguousarray(C.indices, dtype=np.int32)
        data = np.ascontiguousarray(C.data, dtype=dt)
        op = _OrcOp(0, A.shape[0], indptr.ctypes.data, indices.ctypes.data, data.ctypes.data)
        return op, (indptr, indices, data)
    M = np.asfortranarray(np.asarray(A, dtype=dt))
    op = _OrcOp(1, A.shape[0], None, None, M.ctypes.data)
    return op, (M,)


def c_lanczos(A, v: np.ndarray, deg: int, rtol: float = 1e-8, orth: int = 0, ncv: Optional[int] = None, _op=None):
    """C restatement (oracle/slq_oracle.c: orc_lanczos_f64/f32).  Returns (alpha, beta, Q, steps)."""
    dt = np.dtype(v.dtype)
    n = A.shape[0]
    if ncv is None:
        orth, ncv = _ncv_for(deg, orth)
    op, keep = _op if _op is not None else _make_orc_op(A, dt)
    alpha, beta = np.zeros(deg + 1, dtype=dt), np.zeros(deg + 1, dtype=dt)
    Q = np.zeros((n, ncv), dtype=dt, order="F")
    q = np.array(v, dtype=dt, copy=True)
    fn = c_lib().orc_lanczos_f64 if dt == np.float64 else c_lib().orc_lanczos_f32
    steps = fn(ctypes.byref(op), q.ctypes.data, deg, float(rtol), orth, alpha.ctypes.data, beta.ctypes.data, Q.ctypes.data, ncv)
    return alpha, beta, Q, steps


# --------------------------------------------------------------------------------------------------
# quadrature: src/primate/integrate.py:58-64, tridiag.py:7-43, tqli.py:15-90, fttr.py:5-29
# --------------------------------------------------------------------------------------------------
def ref_quadrature(alpha: np.ndarray, beta: np.ndarray, deg: int):
    """Golub-Welsch exactly as the reference's default path: SciPy eigh_tridiagonal (LAPACK stemr) on
    (d[:deg], e[1:deg]) -> theta ascending, tau = Y[0,:]**2   (integrate.py:63-64; tridiag.py:10-11)."""
    from scipy.linalg import eigh_tridiagonal

    d, e = np.asarray(alpha)[:deg], np.asarray(beta)[:deg]
    theta, Y = eigh_tridiagonal(d, e[1:])
    return theta, np.square(Y[0, :])


def np_tqli(d: np.ndarray, e: np.ndarray, with_vectors: bool = True, max_iter: int = 30):
    """C port of tqli.py:15-90 (incl. the non-textbook sign, :5-7).  e has a leading zero.  Returns (w, Z)."""
    d = np.array(d, dtype=np.float64, copy=True)
    e = np.array(e, dtype=np.float64, copy=True)
    n = len(d)
    Z = np.eye(n, dtype=np.float64) if with_vectors else None
    c_lib().orc_tqli(d.ctypes.data, e.ctypes.data, Z.ctypes.data if with_vectors else None, n, max_iter)
    return d, Z


def c_fttr(theta: np.ndarray, alpha: np.ndarray, beta: np.ndarray, k: int) -> np.ndarray:
    theta = np.ascontiguousarray(theta, dtype=np.float64)
    alpha = np.ascontiguousarray(alpha, dtype=np.float64)
    beta = np.ascontiguousarray(beta, dtype=np.float64)
    w = np.zeros(len(theta), dtype=np.float64)
    c_lib().orc_fttr(theta.ctypes.data, alpha.ctypes.data, beta.ctypes.data, len(alpha), k, w.ctypes.data)
    return w


# --------------------------------------------------------------------------------------------------
# MatrixFunction.quad / hutch: src/primate/operators.py:126-151, trace.py:86-116, stats.py:66-86
# --------------------------------------------------------------------------------------------------
def ref_quad(A, X: np.ndarray, fun="identity", deg: int = 20, orth: int = 0, rtol: float = 1e-8, dtype=np.float64,
             return_all: bool = False, lanczos: Callable = ref_lanczos, **fun_kw):  # fmt: skip
    """operators.py:138-151 per column, FRESH Q per probe, ncv = deg (operators.py:77).

    Returns y (k,) float64; with ``return_all`` also per-probe (alpha, beta, theta, tau)."""
    f = spectral_fun(fun, **fun_kw)
    n = A.shape[0]
    deg = min(deg, n)
    orth = deg if orth < 0 or orth > deg else orth  # operators.py:79
    X = np.asarray(X, dtype=dtype)
    X = X[:, None] if X.ndim == 1 else X
    y = np.zeros(X.shape[1])
    extra = []
    # one basis buffer for all columns, as the reference holds `_Q` (operators.py:74,138-148) -- without it every probe
    # would page-fault a fresh n x deg array (400 MB at config C2), a cost the reference does not pay
    kw = {"Q": np.zeros((n, deg), dtype=np.dtype(dtype), order="F")} if lanczos is ref_lanczos else {}
    for j in range(X.shape[1]):
        xc = X[:, j]
        x_norm_sq = np.linalg.norm(xc) ** 2  # operators.py:147
        out = lanczos(A, xc, deg, rtol, orth, ncv=deg, **kw)
        alpha, beta = out[0], out[1]
        theta, tau = ref_quadrature(alpha, beta, deg)  # operators.py:149
        y[j] = np.sum(f(theta) * tau, axis=-1) * x_norm_sq  # operators.py:150
        if return_all:
            extra.append((alpha.copy(), beta.copy(), theta, tau))
    return (y, extra) if return_all else y


def ref_matvec(A, x: np.ndarray, fun="identity", deg: int = 20, orth: int = 0, rtol: float = 1e-8, dtype=np.float64, **fun_kw):
    """operators.py:102-124: y = ||x|| Q Y (f(theta) o Y[0,:])^T with the full basis (ncv = deg), zeroed per call
    (operators.py:116), SciPy eigh_tridiagonal for (theta, Y) (tridiag.py:10-11).  x: (n,) or (n, m) -> same shape."""
    from scipy.linalg import eigh_tridiagonal

    f = spectral_fun(fun, **fun_kw)
    n = A.shape[0]
    deg = min(deg, n)
    orth = deg if orth < 0 or orth > deg else orth
    X = np.asarray(x, dtype=dtype)
    cols = X[:, None] if X.ndim == 1 else X
    out = np.empty(cols.shape, dtype=np.dtype(dtype), order="F")
    Q = np.zeros((n, deg), dtype=np.dtype(dtype), order="F")
    for j in range(cols.shape[1]):
        xc = np.ascontiguousarray(cols[:, j])
        x_norm = np.linalg.norm(xc)
        Q.fill(0)  # operators.py:116
        alpha, beta, _ = ref_lanczos(A, xc, deg, rtol, orth, ncv=deg, Q=Q)
        rw, Y = eigh_tridiagonal(alpha[:deg], beta[1:deg])  # operators.py:122
        out[:, j] = x_norm * (Q @ (Y @ (np.ravel(f(rw)) * Y[0, :])))  # operators.py:124
    return out[:, 0] if X.ndim == 1 else out


class Welford:
    """stats.py:66-86 (Covariance.update, dim=1) + estimators.py:128-141 (delta, n_samples)."""

    def __init__(self):
        self.n, self.mu, self.S, self.delta = 0, 0.0, 0.0, np.inf

    def update(self, x):
        x = np.atleast_1d(np.asarray(x, dtype=np.float64)).ravel()
        m = x.shape[0]
        batch_mean = x.mean()
        delta_mean = batch_mean - self.mu
        new_n = self.n + m
        old_mu = self.mu
        self.mu = self.mu + (m / new_n) * delta_mean
        xc = x - batch_mean
        self.S += float(xc @ xc) + (self.n * m / new_n) * (delta_mean * delta_mean)
        self.n = new_n
        self.delta = self.mu - old_mu

    @property
    def variance(self):
        return self.S / (self.n - 1) if self.n > 1 else np.inf


def ref_hutch(A, fun="identity", deg: int = 20, orth: int = 0, count: int = 200, seed=None, batch: int = 1,
              dtype=np.float64, lanczos: Callable = ref_lanczos, first_probe: int = 0, return_samples: bool = False,
              **fun_kw):  # fmt: skip
    """trace.py:86-116 with converge="count": draw `batch` probes per iteration from ONE PCG64 stream
    (probe i = stream elements [i*N, (i+1)*N): SURVEY A.1), update the running mean, stop at len >= count.

    ``first_probe`` starts the stream at probe index `first_probe` via PCG64.advance (used to shard the
    CPU baseline over processes)."""
    N = A.shape[0]
    rng = np.random.default_rng(seed)
    if first_probe:
        rng.bit_generator.advance(first_probe * N)
    est, samples = Welford(), []
    while est.n < count:
        v = ref_isotropic((N, batch), "rademacher", rng).astype(dtype)  # trace.py:108,115
        y = ref_quad(A, v, fun, deg, orth, dtype=dtype, lanczos=lanczos, **fun_kw)
        est.update(y)
        samples.extend(y.tolist())
    return (est.mu, np.array(samples)) if return_samples else est.mu


# --------------------------------------------------------------------------------------------------
# block drivers: src/primate/trace.py:119-315, diagonal.py:11-138 (host algebra on a `matmat` callable)
# --------------------------------------------------------------------------------------------------
def _coldots(X, Y):
    """diag(X^T Y) as a column (the reference's `diag_prod`, trace.py:201 / diagonal.py:111)."""
    return np.einsum("ij,ij->j", X, Y)[:, None]


def ref_hutchpp(matmat: Callable, n: int, m: int, pdf: str = "rademacher", seed=None, dtype=np.float64, quad: Optional[Callable] = None):
    """trace.py:135-182: sketch Q = qr(A W), exact trace on range(Q) (`quad` per column in "reduced" mode, else
    diag(Q^T A Q)), Hutchinson on the deflated remainder.  Returns (estimate, samples)."""
    rng = np.random.default_rng(seed)
    nb = m + m % 3  # trace.py:153-154
    WB = ref_isotropic((n, nb), pdf, rng).astype(dtype)  # trace.py:159
    Q = np.linalg.qr(matmat(WB), mode="reduced")[0]  # trace.py:160
    rng_ests = np.ravel(quad(Q)) if quad is not None else np.einsum("ij,ij->j", matmat(Q), Q)  # trace.py:168
    G = ref_isotropic((n, nb), pdf, rng).astype(dtype)  # trace.py:172
    G -= Q @ (Q.T @ G)  # trace.py:173
    defl_ests = np.einsum("ij,ij->j", matmat(G), G)  # trace.py:174
    est = np.sum(rng_ests) + np.sum(defl_ests) / nb  # trace.py:169,175-178
    return est, np.concatenate([np.ravel(rng_ests), defl_ests])


def ref_xdiag(matmat: Callable, n: int, m: Optional[int] = None, pdf: str = "sphere", seed=None):
    """diagonal.py:99-138 (XDiag) on a symmetric operator (A.T @ Q == A @ Q)."""
    m = 2 * n if m is None else min(m + (m % 2), 2 * n)  # diagonal.py:107
    m = m // 2
    rng = np.random.default_rng(seed=seed)
    N = ref_isotropic((n, m), pdf, rng)  # diagonal.py:115
    Y = matmat(N)
    Q, R = np.linalg.qr(Y, mode="reduced")  # diagonal.py:117
    rowdot = lambda X, Z: np.einsum("ij,ij->i", X, Z)[:, None]  # noqa: E731  (diag_prod(X.T, Z.T), diagonal.py:111)
    dNY = rowdot(N, Y)
    Z = matmat(Q)  # diagonal.py:123
    T = Z.T @ N
    R_inv = np.linalg.inv(R)
    S = R_inv.T / np.linalg.norm(R_inv, axis=1)  # diagonal.py:126
    QS = Q @ S
    dQZ, dQSSZ, dNTQ = rowdot(Q, Z), rowdot(QS, Z @ S), rowdot(N, Q @ T)  # diagonal.py:130-132
    dNQSST = rowdot(N, QS * _coldots(S, T).T)  # diagonal.py:133
    return (dQZ + (-dQSSZ + dNY - dNTQ + dNQSST) / m).ravel()  # diagonal.py:137


def _ref_xtrace_samples(W, Z, Q, R, R_inv):
    """trace.py:185-227 with the non-"sphere" scale (xtrace hands `_xtrace` the sampler callable, so the sphere branch
    at trace.py:209-214 is never taken: SURVEY A.2)."""
    m = W.shape[1]
    Wp = Q.T @ W
    S = R_inv.T / np.linalg.norm(R_inv, axis=1)  # trace.py:206
    H = Q.T @ Z
    HW = H @ Wp
    T = Z.T @ W
    dSW, dSHS = _coldots(S, Wp), _coldots(S, H @ S)
    dTW, dWHW = _coldots(T, Wp), _coldots(Wp, HW)
    dSRmHW, dTmHRS = _coldots(S, R - HW), _coldots(T - H.T @ Wp, S)
    ests = np.trace(H) * np.ones((m, 1)) - dSHS  # trace.py:224
    ests += -dTW + dWHW + dSW * dSRmHW + np.abs(dSW) ** 2 * dSHS + dTmHRS * dSW  # trace.py:225
    return ests


def ref_xtrace(matmat: Callable, n: int, batch: int = 32, pdf: str = "sphere", seed=None, max_samples: Optional[int] = None):
    """trace.py:268-315: grows W, Q, R, Z batch by batch until n samples (the reference's CountCriterion(n) is always
    in force, trace.py:271-276); `max_samples` truncates the loop for tests.  Returns (estimate, per-batch estimates)."""
    rng = np.random.default_rng(seed)
    stop = n if max_samples is None else min(n, max_samples)
    W, Ymat, Z = np.zeros((n, 0)), np.zeros((n, 0)), np.zeros((n, 0))
    history, est = [], np.nan
    while W.shape[1] < stop:
        ns = min(n - W.shape[1], int(batch))  # trace.py:289
        Nw = ref_isotropic((n, ns), pdf, rng)
        Ymat = np.c_[Ymat, matmat(Nw)]
        Q, R = np.linalg.qr(Ymat, mode="reduced")  # == the qr_insert sequence up to column signs, which cancel in _xtrace
        R_inv = np.linalg.inv(R)
        W = np.c_[W, Nw]
        Z = np.c_[Z, matmat(Q[:, -ns:])]  # trace.py:303 -- NB images of the columns as they were when appended
        # qr_insert keeps earlier columns of Q fixed, so Z's earlier columns stay valid (QR of a prefix is a prefix)
        est = float(np.mean(_ref_xtrace_samples(W, Z, Q, R, R_inv)))
        history.append(est)
    return est, history


def ref_diag(matmat: Callable, n: int, count: int, pdf: str = "rademacher", seed=None, dtype=np.float64):
    """diagonal.py:84-91 with converge="count": running ratio sum(v * Av) / sum(v * v), averaged per probe by a
    MeanEstimator(dim=n) (estimators.py:128-141)."""
    rng = np.random.default_rng(seed)
    numer, denom, mu = np.zeros(n, dtype=dtype), np.zeros(n, dtype=dtype), np.zeros(n)
    for i in range(count):
        v = ref_isotropic((n, 1), pdf, rng).astype(dtype)
        u = np.ravel(matmat(v))
        numer += u * v.ravel()
        denom += np.square(v.ravel())
        mu += ((numer / denom) - mu) / (i + 1)
    return mu
