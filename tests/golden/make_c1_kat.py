"""Generates tests/golden/c1_kat.json: known answers for BASELINE config C1 -- the literal 6x6 system of the
reference example (examples/pcg_solve.cu:14-25, same numbers in examples/pcg_solve_dp.cu:14-25).

The reference ships no expected output, so the answers are DERIVED here, independently of
oracle/pcg_oracle.c and of the CUDA code, with dense numpy:
  * lambda_star = S^-1 gamma by fp64 LU,
  * the fp64 / fp32 eta sequences, iteration counts and solutions of the recurrence of reference
    include/pcg.cuh:118-212 written with dense matrices,
for Phi^-1 in {identity, block-Jacobi, symmetric stair}.   Run:  python tests/golden/make_c1_kat.py
"""
import json
import os

import numpy as np

S_FLAT = [0, 0, 0, 0, -.999, 0, 0, -.999, .999, .0999, -.98, .999,
          .999, -.98, .0999, .999, -2.008, .8801, .8801, -3.0584, .999, .0999, -.98, .999,
          .999, -.98, .0999, .999, -1.019, .8801, .8801, -2.0694, 0, 0, 0, 0]
GAMMA = [3.1385, 0, 0, 3.0788, .0031, 3.0788]
NX, N = 2, 3


def dense(flat):
    M = np.asarray(flat, dtype=np.float64).reshape(N, 3, NX, NX)   # [k, j, c, r]: column-major blocks
    A = np.zeros((N * NX, N * NX))
    for k in range(N):
        for j in range(3):
            kc = k + j - 1
            if 0 <= kc < N:
                A[k * NX:(k + 1) * NX, kc * NX:(kc + 1) * NX] = M[k, j].T
    return A


def precond(A, kind):
    n = N * NX
    P = np.zeros((n, n))
    blk = lambda i, j: A[i * NX:(i + 1) * NX, j * NX:(j + 1) * NX]
    if kind == "identity":
        return np.eye(n)
    Dinv = [np.linalg.inv(blk(k, k)) for k in range(N)]
    for k in range(N):
        P[k * NX:(k + 1) * NX, k * NX:(k + 1) * NX] = Dinv[k]
    if kind == "stair":
        for k in range(N - 1):
            U = -Dinv[k] @ blk(k, k + 1) @ Dinv[k + 1]
            P[k * NX:(k + 1) * NX, (k + 1) * NX:(k + 2) * NX] = U
            P[(k + 1) * NX:(k + 2) * NX, k * NX:(k + 1) * NX] = U.T
    return P


def pcg(A, P, g, dt, tol=1e-6, max_iter=25):
    A, P, g = A.astype(dt), P.astype(dt), g.astype(dt)
    lam = np.zeros_like(g)
    r = g - A @ lam
    rt = P @ r
    p = rt.copy()
    eta = dt(r @ rt)
    etas = [float(eta)]
    it, hit = 0, True
    while it < max_iter:
        ups = A @ p
        alpha = dt(eta / dt(p @ ups))
        lam = lam + alpha * p
        r = r - alpha * ups
        rt = P @ r
        eta_new = dt(r @ rt)
        etas.append(float(eta_new))
        it += 1
        if abs(eta_new) < tol:
            hit = False
            break
        beta = dt(eta_new / eta)
        eta = eta_new
        p = rt + beta * p
    return lam, it, hit, etas


def main():
    A = dense(S_FLAT)
    g = np.asarray(GAMMA)
    out = {"source": "derived with dense numpy from reference examples/pcg_solve.cu:14-25 (no expected output upstream)",
           "nx": NX, "N": N, "S": S_FLAT, "gamma": GAMMA, "exit_tol": 1e-6, "max_iter": 25,
           "lambda_star": np.linalg.solve(A, g).tolist(), "eigenvalues": np.linalg.eigvalsh(A).tolist(), "cases": []}
    for kind in ("identity", "jacobi", "stair"):
        P = precond(A, kind)
        for dt in (np.float64, np.float32):
            lam, it, hit, etas = pcg(A, P, g, dt)
            out["cases"].append({"precond": kind, "dtype": np.dtype(dt).name, "iters": it, "max_iter_exit": hit,
                                 "eta": etas, "lambda": [float(x) for x in lam]})
    # what the reference's host overload yields on zero-filled device memory (defect D1): Phi^-1 = 0
    lam, it, hit, etas = pcg(A, np.zeros_like(A), g, np.float64)
    out["cases"].append({"precond": "zero", "dtype": "float64", "iters": it, "max_iter_exit": hit, "eta": None, "lambda": None,
                         "all_nan": bool(np.isnan(lam).all())})
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "c1_kat.json"), "w") as f:
        json.dump(out, f, indent=1)
    for c in out["cases"]:
        print(c["precond"], c["dtype"], c["iters"], c["max_iter_exit"])


if __name__ == "__main__":
    main()
