"""Test helper: the reduced pose system [H | b] as the GPU accumulate stage builds it (fp64, packed lower triangle behind
db200_ba_system_ptr) and as the CPU oracle builds it (dump_system).  Comparing the SYSTEMS is the sharp parity check of
the accumulate stage: a wrong or missing block shows as >= 1e-3, fp32 round-off as ~1e-6 — whereas dx amplifies that
round-off by the condition number of the damped system (1e5 .. 1e6 on backend graphs with small maps)."""
import ctypes as C

import numpy as np
import torch


def gpu_system(p):
    from droid_b200 import _capi
    lib = _capi.lib()
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    N, (B, ht, wd), M = p.n_edges, p.disps.shape, p.eta.shape[0]
    wsb = lib.db200_ba_workspace_bytes(N, B, ht, wd, p.t0, p.t1, M, 0)
    ws = torch.zeros(wsb, dtype=torch.uint8, device="cuda")
    g = [t(x) for x in (p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, p.jj)]
    P_ = lambda x: C.c_void_p(x.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.db200_ba_accumulate(*[P_(x) for x in g], N, B, ht, wd, p.t0, p.t1, 0, M, 1, 0, B, P_(ws), wsb, st)
    assert rc == 0 and lib.db200_ba_status(P_(ws), st) == 0
    n = 6 * (p.t1 - p.t0)
    off = lib.db200_ba_system_ptr(P_(ws), N, B, ht, wd, p.t0, p.t1, M, 0) - ws.data_ptr()
    hb = ws[off:off + 8 * (n * (n + 1) // 2 + n)].view(torch.float64).cpu().numpy()
    H = np.zeros((n, n))
    H[np.tril_indices(n)] = hb[:n * (n + 1) // 2]
    return H + np.tril(H, -1).T, hb[n * (n + 1) // 2:].copy()


def oracle_system(p, lm=1e-5, ep=1e-2):
    from oracle import oracle
    r = oracle.ba(p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, p.jj, p.t0, p.t1, 1, lm, ep,
                  dump_system=True)
    H = r["H"]
    return np.tril(H) + np.tril(H, -1).T, r["b"]


def check_system(p, tol=2e-5):
    """normwise, and per 6x6 block row so that a single wrong block cannot hide in the norm of the whole matrix"""
    H, b = gpu_system(p)
    Ho, bo = oracle_system(p)
    rel = lambda x, y: float(np.linalg.norm(x - y) / max(np.linalg.norm(y), 1e-300))
    eh, eb = rel(H, Ho), rel(b, bo)
    n = H.shape[0]
    rows = max(rel(H[r:r + 6], Ho[r:r + 6]) for r in range(0, n, 6))
    assert eh <= tol and eb <= tol and rows <= 5 * tol, (eh, eb, rows)
    assert ((Ho != 0) == (H != 0)).mean() > 0.999          # same block-sparsity pattern
    return eh, eb, rows


def check_solve(p, lm, ep, tol=1e-5):
    """the solve stage alone: the GPU's dx against a numpy fp64 solve of the GPU's OWN system with the reference's damping
    diag += ep + lm * diag (droid_kernels.cu:1197).  With check_system this splits `dx` parity into its two factors:
    system at fp32 round-off x exact solve; what remains between two such dx is the conditioning of the problem."""
    import droid_backends
    H, b = gpu_system(p)
    n = H.shape[0]
    A = H.copy()
    dg = np.diag(A).copy()
    A[np.diag_indices(n)] = dg + ep + lm * dg
    want = np.linalg.solve(A, b).reshape(-1, 6)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    dx, _ = droid_backends.ba(t(p.poses), t(p.disps), t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta), t(p.ii),
                              t(p.jj), p.t0, p.t1, 1, lm, ep, False)
    e = float(np.linalg.norm(dx.cpu().numpy() - want) / np.linalg.norm(want))
    assert e <= tol, e
    return e
