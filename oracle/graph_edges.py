"""ORACLE (test infrastructure only — never imported by the product path).

CPU restatement of the reference's edge selection, line by line in plain Python/numpy:
  * `add_proximity_factors`  /root/reference/droid_slam/factor_graph.py:315-379
  * `__filter_repeated_edges` /root/reference/droid_slam/factor_graph.py:44-55
  * `add_neighborhood_factors` /root/reference/droid_slam/factor_graph.py:302-312
The pairwise distances `d` come from `oracle.frame_distance` (pinned bit-exactly against the reference's CUDA
kernel by tests/golden), averaged as depth_video.py:171-181 does.  Pinning of THIS file: the reference's three methods
are run UNMODIFIED (factor_graph.py imported under the lietorch shim, with a stub `video` whose `distance()` is that
average) by tests/golden/make_pyref_graph_golden.py on the 8 cases of tests/graph_cases.py; the lists are committed as
tests/golden/pyref_graph_edges.npz and tests/test_oracle.py asserts this restatement reproduces them element for element
(tests/test_gpu_graph.py asserts the same of db200_proximity_edges).
`torch.argsort` leaves the order of equal keys unspecified; a stable sort is used (lower flat index first).
"""
import numpy as np


def proximity_distance(frame_distance, poses, disps, intrinsics, t0, t1, t, beta):
    """depth_video.py:160-190 on meshgrid(arange(t0,t), arange(t1,t)) (factor_graph.py:318-326)."""
    ix, jx = np.arange(t0, t), np.arange(t1, t)
    ii, jj = np.meshgrid(ix, jx, indexing="ij")
    ii, jj = ii.reshape(-1).astype(np.int64), jj.reshape(-1).astype(np.int64)
    if ii.size == 0:
        return ii, jj, np.zeros(0, np.float32)
    p = np.ascontiguousarray(poses[:t])
    d1 = frame_distance(p, disps, intrinsics, ii, jj, beta)
    d2 = frame_distance(p, disps, intrinsics, jj, ii, beta)
    return ii, jj, (np.float32(0.5) * (d1.astype(np.float32) + d2.astype(np.float32))).astype(np.float32)


def add_proximity_factors_edges(ii, jj, d, t0, t1, t, rad, nms, thresh, max_factors, stereo, ii1, jj1):
    """factor_graph.py:327-379 from the distances on; returns the list `es` of (i, j) tuples."""
    d = np.array(d, dtype=np.float32, copy=True)
    if d.size:
        d[ii - rad < jj] = np.inf
        d[d > 100] = np.inf
    cols = t - t1

    def suppress(i, j):
        for di in range(-nms, nms + 1):
            for dj in range(-nms, nms + 1):
                if abs(di) + abs(dj) <= max(min(abs(i - j) - 2, nms), 0):
                    i1, j1 = i + di, j + dj
                    if (t0 <= i1 < t) and (t1 <= j1 < t):
                        d[(i1 - t0) * cols + (j1 - t1)] = np.inf

    for i, j in zip(np.asarray(ii1).tolist(), np.asarray(jj1).tolist()):
        suppress(i, j)

    es = []
    for i in range(t0, t):
        if stereo:
            es.append((i, i))
            d[(i - t0) * cols + (i - t1)] = np.inf            # numpy wraps a negative index like torch does
        for j in range(max(i - rad - 1, 0), i):
            es.append((i, j))
            es.append((j, i))
            d[(i - t0) * cols + (j - t1)] = np.inf

    order = np.argsort(d, kind="stable")
    for k in order:
        if float(d[k]) > thresh:
            continue
        if len(es) > max_factors:
            break
        i, j = int(ii[k]), int(jj[k])
        es.append((i, j))
        es.append((j, i))
        suppress(i, j)
    return es


def filter_repeated_edges(ii, jj, ii_old, jj_old):
    eset = set(zip(np.asarray(ii_old).tolist(), np.asarray(jj_old).tolist()))
    keep = np.array([(i, j) not in eset for i, j in zip(np.asarray(ii).tolist(), np.asarray(jj).tolist())], dtype=bool)
    return np.asarray(ii)[keep], np.asarray(jj)[keep]


def add_neighborhood_factors_edges(t0, t1, r=3, stereo=False):
    ii, jj = np.meshgrid(np.arange(t0, t1), np.arange(t0, t1), indexing="ij")
    ii, jj = ii.reshape(-1), jj.reshape(-1)
    c = 1 if stereo else 0
    keep = (np.abs(ii - jj) > c) & (np.abs(ii - jj) <= r)
    return ii[keep].astype(np.int64), jj[keep].astype(np.int64)
