"""Generates tests/golden/slq_golden.npz by importing the UNMODIFIED reference package.

Runs only in the build container (needs /root/reference and oracle/_ref/_lanczos*.so); the GPU box and
the test-suite only ever read the committed ``.npz``.  The reference is imported without copying any of
its sources: a namespace-style package object ``primate`` is created whose ``__path__`` lists the
reference's ``src/primate`` directory (pure-Python modules) and ``oracle/_ref`` (the compiled
``_lanczos`` extension built by ``oracle/build_oracle.py``).  The reference's ``__init__`` is skipped
because it asks importlib.metadata for an installed distribution (``__init__.py:3``).

    python tests/golden/make_golden.py
"""
import os
import sys
import types

import numpy as np
import scipy.sparse as sp

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF_SRC = "/root/reference/src/primate"
sys.path.insert(0, ROOT)


def import_reference():
    from oracle.build_oracle import build_reference_extension

    build_reference_extension(verbose=False)
    pkg = types.ModuleType("primate")
    pkg.__path__ = [REF_SRC, os.path.join(ROOT, "oracle", "_ref")]
    pkg.__version__ = "0.5.6"
    sys.modules["primate"] = pkg
    import primate.lanczos  # noqa: F401
    import primate.operators  # noqa: F401
    import primate.trace  # noqa: F401
    import primate.diagonal  # noqa: F401
    import primate.integrate  # noqa: F401
    import primate.random  # noqa: F401
    import primate.tridiag  # noqa: F401
    import primate.tqli  # noqa: F401
    import primate.fttr  # noqa: F401

    return pkg


def laplacian_2d(g: int, dtype=np.float64):
    T = sp.diags([-1.0, 2.0, -1.0], [-1, 0, 1], shape=(g, g), dtype=dtype)
    I = sp.identity(g, dtype=dtype)  # noqa: E741
    return (sp.kron(I, T) + sp.kron(T, I)).tocsr().astype(dtype)


def fresh_quad(MF, A, fun, deg, orth, V, **kw):
    """`MatrixFunction.quad` per column on a FRESH instance (zeroed _Q, _alpha, _beta): SURVEY fact 4."""
    out, alphas, betas, nodes, weights = [], [], [], [], []
    for j in range(V.shape[1]):
        M = MF(A, fun=fun, deg=deg, orth=orth, **kw)
        out.append(M.quad(V[:, [j]]).item())
        alphas.append(M._alpha.copy())
        betas.append(M._beta.copy())
        nodes.append(M._nodes.copy())
        weights.append(M._weights.copy())
    return np.array(out), np.array(alphas), np.array(betas), np.array(nodes), np.array(weights)


def main():
    import_reference()
    from primate.diagonal import diag, xdiag
    from primate.fttr import fttr
    from primate.integrate import quadrature
    from primate.lanczos import lanczos
    from primate.operators import MatrixFunction
    from primate.random import isotropic, symmetric
    from primate.tqli import tqli
    from primate.trace import hutch, hutchpp, xtrace
    from primate.tridiag import eigh_tridiag

    G = {}
    # ---- dense SPD fixture from the reference docs (SURVEY 8c goldens) -------------------------------
    A = symmetric(150, pd=True, seed=1234)
    G["A150"] = A
    G["A150_trace"] = A.trace()
    G["A150_logdet"] = np.sum(np.log(np.linalg.eigvalsh(A)))

    a, b = lanczos(A, deg=10, seed=0, orth=0)
    G["lz_seed0_deg10_a"], G["lz_seed0_deg10_b"] = a.copy(), b.copy()

    v = isotropic((150, 1), "rademacher", seed=42)
    G["v42"] = v.ravel().copy()
    for orth in (0, 3, 20):
        (a, b), Q = lanczos(A, v0=v.ravel(), deg=20, orth=orth, return_basis=True)
        G[f"lz_v42_deg20_orth{orth}_a"], G[f"lz_v42_deg20_orth{orth}_b"] = a.copy(), b.copy()
        G[f"lz_v42_deg20_orth{orth}_Q"] = np.array(Q)
    # ncv = clip(orth, 2, deg) path (no basis)
    a, b = lanczos(A, v0=v.ravel(), deg=20, orth=5)
    G["lz_v42_deg20_orth5_nobasis_a"], G["lz_v42_deg20_orth5_nobasis_b"] = a.copy(), b.copy()

    # fp32
    A32 = A.astype(np.float32)
    a, b = lanczos(A32, v0=v.ravel().astype(np.float32), deg=20, orth=0)
    G["lz32_v42_deg20_a"], G["lz32_v42_deg20_b"] = a.copy(), b.copy()

    # ---- quad for each builtin --------------------------------------------------------------------
    V4 = isotropic((150, 4), "rademacher", seed=7)
    G["V4_seed7"] = np.array(V4)
    funs = {
        "identity": {}, "abs": {}, "sqrt": {}, "log": {}, "inv": {}, "exp": {"t": -0.5},
        "smoothstep": {"a": 0.2, "b": 0.6}, "numrank": {"threshold": 0.3},
    }  # fmt: skip
    for fun, kw in funs.items():
        for orth in (0, 3):
            y, al, be, nd, wt = fresh_quad(MatrixFunction, A, fun, 20, orth, V4, **kw)
            G[f"quad_{fun}_orth{orth}"] = y
            if fun == "log":
                G[f"quad_log_orth{orth}_alpha"], G[f"quad_log_orth{orth}_beta"] = al, be
                G[f"quad_log_orth{orth}_nodes"], G[f"quad_log_orth{orth}_weights"] = nd, wt
    M = MatrixFunction(A, "log", deg=20, orth=0)
    G["quad_log_v42"] = M.quad(v).item()

    # ---- hutch ------------------------------------------------------------------------------------
    M = MatrixFunction(A, "log", deg=20, orth=0)
    G["hutch_log_count200_seed5"] = hutch(M, converge="count", count=200, seed=5)
    est, res = hutch(M, converge="count", count=200, seed=5, full=True, batch=32, record=True)
    G["hutch_log_count200_seed5_full_b32"] = est
    G["hutch_log_count200_seed5_full_b32_samples"] = np.array(res.estimator.values).ravel()
    est, res = hutch(M, seed=11, full=True, batch=1, record=True)  # default: Count(200) | Confidence(.95, atol=1)
    G["hutch_log_default_seed11"] = est
    G["hutch_log_default_seed11_n"] = len(res.estimator)
    est, res = hutch(M, converge="confidence", confidence=0.9, atol=2.0, rtol=0.0, seed=3, full=True, batch=8, record=True)
    G["hutch_log_conf_seed3_b8"] = est
    G["hutch_log_conf_seed3_b8_n"] = len(res.estimator)
    G["hutch_A_default_seed2"] = hutch(A, seed=2)  # plain matrix (einsum branch, trace.py:97)
    M3 = MatrixFunction(A, "exp", deg=12, orth=0, t=0.25)
    G["hutch_exp_count64_seed9"] = hutch(M3, converge="count", count=64, seed=9)

    # ---- f(A) v  (operators.py:102-124) ---------------------------------------------------------------
    for fun, kw in (("log", {}), ("exp", {"t": -0.5}), ("identity", {})):
        for orth in (0, 20):
            M = MatrixFunction(A, fun, deg=20, orth=orth, **kw)
            G[f"matvec_{fun}_orth{orth}_v42"] = np.asarray(M._matvec(v.ravel().copy())).ravel()

    # ---- sparse 2-D Laplacian (config-2 shape, small) ------------------------------------------------
    L = laplacian_2d(30)
    G["L30_indptr"], G["L30_indices"], G["L30_data"] = L.indptr.copy(), L.indices.copy(), L.data.copy()
    VL = isotropic((900, 3), "rademacher", seed=0)
    G["VL_seed0"] = np.array(VL)
    for orth in (0, 3):
        y, al, be, nd, wt = fresh_quad(MatrixFunction, L, "log", 50, orth, VL)
        G[f"L30_quad_log_orth{orth}"] = y
        G[f"L30_quad_log_orth{orth}_alpha"], G[f"L30_quad_log_orth{orth}_beta"] = al, be
    ML = MatrixFunction(L, "log", deg=50, orth=0)
    G["L30_hutch_log_count64_seed0"] = hutch(ML, converge="count", count=64, seed=0)
    L32 = L.astype(np.float32)
    a, b = lanczos(L32, v0=VL[:, 0].astype(np.float32), deg=50, orth=0)
    G["L30_lz32_a"], G["L30_lz32_b"] = a.copy(), b.copy()

    # ---- early break: operator with 3 distinct eigenvalues -> invariant subspace at step 3 -----------
    rng = np.random.default_rng(3)
    Qh, _ = np.linalg.qr(rng.standard_normal((60, 60)))
    ew = np.repeat([0.5, 1.5, 4.0], 20)
    B = (Qh * ew) @ Qh.T
    B = (B + B.T) / 2
    G["B60"] = B
    vb = isotropic((60, 2), "rademacher", seed=1)
    G["VB_seed1"] = np.array(vb)
    for j in range(2):
        al, be = np.zeros(11), np.zeros(11)
        lanczos(B, v0=vb[:, j], deg=10, orth=0, alpha=al, beta=be)
        G[f"B60_lz_alpha_{j}"], G[f"B60_lz_beta_{j}"] = al, be
    y, al, be, nd, wt = fresh_quad(MatrixFunction, B, "log", 10, 0, vb)
    G["B60_quad_log"] = y
    y, _, _, _, _ = fresh_quad(MatrixFunction, B, "exp", 10, 0, vb)
    G["B60_quad_exp"] = y

    # ---- tridiagonal solvers / quadrature -----------------------------------------------------------
    rng = np.random.default_rng(1234)
    d = rng.uniform(0.5, 2.0, size=40)
    e = np.append([0.0], rng.uniform(0.1, 1.0, size=39))
    G["T40_d"], G["T40_e"] = d, e
    nodes, weights = quadrature(d, e, deg=40, quad="gw")
    G["T40_gw_nodes"], G["T40_gw_weights"] = nodes.copy(), weights.copy()
    nodes, weights = quadrature(d, e, deg=25, quad="gw")
    G["T40_gw25_nodes"], G["T40_gw25_weights"] = nodes.copy(), weights.copy()
    nodes_f, weights_f = quadrature(d, e, deg=40, quad="fttr")
    G["T40_fttr_nodes"], G["T40_fttr_weights"] = nodes_f.copy(), weights_f.copy()
    w = np.zeros(40)
    fttr(np.array(G["T40_gw_nodes"]), d, e, 40, w)
    G["T40_fttr_direct"] = w
    dd, ee, Z = d.copy(), e.copy(), np.eye(40)
    tqli(dd, ee, Z, 30)
    G["T40_tqli_w"], G["T40_tqli_Z"] = dd, Z
    rw, Y = eigh_tridiag(d, e)
    G["T40_eigh_w"], G["T40_eigh_Y0sq"] = rw, Y[0, :] ** 2

    # ---- raw probe streams (bit-exact gate) ----------------------------------------------------------
    G["raw_seed0_first64"] = np.random.PCG64(0).random_raw(64)
    W = isotropic((1000, 5), "rademacher", seed=123)
    G["signs_seed123_1000x5"] = (np.array(W) > 0).astype(np.uint8)
    W = isotropic((257, 33), "rademacher", seed=99)
    G["signs_seed99_257x33"] = (np.array(W) > 0).astype(np.uint8)

    # ---- block drivers on the same engine (API parity, later rows of SURVEY 8f) ----------------------
    G["hutchpp_A_m30_seed4"] = hutchpp(A, m=30, seed=4)
    M = MatrixFunction(A, "exp", deg=20, orth=3, t=0.5)
    G["xdiag_A_m20_seed6"] = xdiag(A, m=20, seed=6)
    G["xtrace_A_b10_seed8"] = xtrace(A, batch=10, seed=8)
    G["diag_A_seed10"] = diag(A, converge="count", count=50, seed=10)

    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "slq_golden.npz")
    np.savez_compressed(out, **{k: np.asarray(v) for k, v in G.items()})
    print(f"wrote {out}: {len(G)} arrays, {os.path.getsize(out) / 1024:.0f} KiB")
    for k in ("A150_trace", "A150_logdet", "quad_log_v42", "hutch_log_count200_seed5", "hutch_log_count200_seed5_full_b32",
              "hutch_log_default_seed11_n", "hutch_log_conf_seed3_b8_n"):  # fmt: skip
        print(k, repr(np.asarray(G[k]).item()))
    print("lz_v42 a[:3]", G["lz_v42_deg20_orth0_a"][:3], "b[:3]", G["lz_v42_deg20_orth0_b"][:3])
    print("B60 beta", G["B60_lz_beta_0"])


if __name__ == "__main__":
    main()
