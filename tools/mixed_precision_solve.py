"""Would a single-precision factorisation + fp64 iterative refinement solve the damped reduced pose system?  (VERDICT r1, item 4c:
"fp32/TF32 factor on tcgen05 + fp64 residual refinement, gated on the 1e-4 parity".)  CPU experiment on the oracle's system:
A = H + diag(ep + lm * diag H), A x = b; fp32 Cholesky of A (plain and Jacobi-scaled), then x <- x + L^-T L^-1 (b - A x) with
the residual in fp64, until the distance to the fp64 solution stops falling.       python tools/mixed_precision_solve.py [c3|c2|c1]"""
import os
import sys
import time

import numpy as np
import scipy.linalg as sl

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [REPO, os.path.join(REPO, "droid-slam_reserch_b200")]
import bench  # noqa: E402
from droid_b200 import synth  # noqa: E402
from oracle import oracle  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "c3"
p = {"c1": synth.config_c1, "c2": synth.config_c2}[wl]() if wl in ("c1", "c2") else bench.make_problem(wl)
lm, ep = (1e-4, 0.1) if wl in ("c1", "c2") else (1e-5, 1e-2)
t = time.time()
r = oracle.ba(p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, p.jj, p.t0, p.t1, 1, lm, ep, dump_system=True)
H = np.tril(r["H"]) + np.tril(r["H"], -1).T
b = r["b"].copy()
n = H.shape[0]
A = H.copy()
dg = np.diag(A).copy()
A[np.diag_indices(n)] = dg + ep + lm * dg
print(f"{wl}: n = {n}, system from the oracle in {time.time() - t:.1f} s")
w = np.linalg.eigvalsh(A)
print(f"cond(A) = {w[-1] / w[0]:.3e}   (lambda_min {w[0]:.3e}, lambda_max {w[-1]:.3e})")
d = 1.0 / np.sqrt(np.diag(A))
As = A * d[:, None] * d[None, :]
ws_ = np.linalg.eigvalsh(As)
print(f"cond(D^-1/2 A D^-1/2) = {ws_[-1] / ws_[0]:.3e}")
x_ref = sl.cho_solve(sl.cho_factor(A, lower=True), b)


def refine(name, M, rhs_scale, unscale, dtype):
    """M = matrix that is factored in `dtype`; correction = unscale(solve(M, rhs_scale(residual)))"""
    try:
        c = sl.cho_factor(M.astype(dtype), lower=True, check_finite=False)
        if not np.all(np.isfinite(c[0])):
            raise np.linalg.LinAlgError("non-finite factor")
    except np.linalg.LinAlgError as e:
        print(f"{name}: factorisation FAILED ({e})")
        return
    x = np.zeros(n)
    hist = []
    for it in range(8):
        res = b - A @ x                                              # fp64 residual
        dx = sl.cho_solve(c, rhs_scale(res).astype(dtype), check_finite=False).astype(np.float64)
        x = x + unscale(dx)
        hist.append(np.linalg.norm(x - x_ref) / np.linalg.norm(x_ref))
    print(f"{name}: |x - x_fp64| / |x_fp64| after 1..8 solves: " + " ".join(f"{h:.1e}" for h in hist))


refine("fp32 Cholesky of A            ", A, lambda v: v, lambda v: v, np.float32)
refine("fp32 Cholesky of the scaled A ", As, lambda v: v * d, lambda v: v * d, np.float32)
# TF32 inputs (10-bit mantissa) with fp32 accumulation, as one tcgen05 kind::tf32 pass would see the operands
tf = lambda M: (M.astype(np.float32).view(np.uint32) + 0x1000 & 0xffffe000).view(np.float32)
refine("fp32 Cholesky of tf32(scaled A)", tf(As).astype(np.float64), lambda v: v * d, lambda v: v * d, np.float32)
