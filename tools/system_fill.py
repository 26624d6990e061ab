"""Tile-level sparsity of the reduced pose system S (6 x 6 blocks per pose) and of its Cholesky factor: how many of the
solver's T x T tiles are structurally non-zero before and after symbolic factorisation, in natural order and after a
reverse Cuthill-McKee ordering of the pose graph.  CPU only.     python tools/system_fill.py [c3|c4] [tile]"""
import os
import sys

import numpy as np
from scipy.sparse import csr_matrix
from scipy.sparse.csgraph import reverse_cuthill_mckee

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [REPO, os.path.join(REPO, "droid-slam_reserch_b200")]
import bench  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "c3"
T = int(sys.argv[2]) if len(sys.argv) > 2 else 64
p = bench.make_problem(wl)
t0, t1 = p.t0, p.t1
P = t1 - t0
A = np.zeros((P, P), bool)
for i in np.unique(p.ii):
    js = np.unique(p.jj[(p.ii == i) & (p.jj != i)])
    grp = np.concatenate([js, [i]]) - t0
    grp = grp[(grp >= 0) & (grp < P)]
    A[np.ix_(grp, grp)] = True
print(f"{wl}: {P} poses, {len(p.ii)} edges; pose-block pattern of S: {A.sum() / P / P:.3f} of the square")


def tiles(perm, name):
    B = A[np.ix_(perm, perm)]
    n = 6 * P
    nt = (n + T - 1) // T
    tile_of = (6 * np.arange(P)) // T, (6 * np.arange(P) + 5) // T             # a pose block can straddle two tiles
    Tm = np.zeros((nt, nt), bool)
    r, c = np.nonzero(B)
    for a in (0, 1):
        for b in (0, 1):
            Tm[tile_of[a][r], tile_of[b][c]] = True
    Tm |= Tm.T
    before = np.tril(Tm).sum()
    L = np.tril(Tm).copy()
    flops_dense = flops = 0
    for k in range(nt):                                       # symbolic right-looking factorisation on tiles
        rows = np.nonzero(L[k + 1:, k])[0] + k + 1
        for a_ in rows:
            L[a_, rows[rows <= a_]] = True
        m = len(rows)
        flops += m * (m + 1) // 2 + m + 1
        md = nt - k - 1
        flops_dense += md * (md + 1) // 2 + md + 1
    tot = nt * (nt + 1) // 2
    # longest chain of dependent panel steps = height of the elimination tree on tiles
    parent = np.full(nt, -1)
    for k in range(nt):
        rows = np.nonzero(L[k + 1:, k])[0]
        if len(rows):
            parent[k] = rows[0] + k + 1
    depth = np.zeros(nt, int)
    for k in range(nt):
        if parent[k] >= 0:
            depth[parent[k]] = max(depth[parent[k]], depth[k] + 1)
    print(f"  {name:14s} tiles {T}: S {before}/{tot} = {before / tot:.3f}, factor {L.sum()}/{tot} = {L.sum() / tot:.3f}, "
          f"tile-GEMM count {flops}/{flops_dense} = {flops / flops_dense:.3f}, elimination-tree height {depth.max() + 1} of {nt}")


tiles(np.arange(P), "natural order")
perm = reverse_cuthill_mckee(csr_matrix(A), symmetric_mode=True)
tiles(np.asarray(perm), "RCM")
