"""Edge-selection cases shared by tests/test_gpu_graph.py, tests/test_oracle.py and tests/golden/make_pyref_graph_golden.py."""
import numpy as np

CASES = [
    # n_kf, t0, t1, rad, nms, beta, thresh, max_factors, stereo, n_old
    (40, 0, 0, 2, 2, 0.25, 16.0, 400, False, 0),          # global backend call (droid_backend.py)
    (40, 0, 0, 2, 2, 0.25, 16.0, 400, True, 30),          # stereo + existing edges
    (40, 35, 15, 2, 2, 0.25, 16.0, 96, False, 20),        # frontend call: t0 = t-5, t1 = t-25 (droid_frontend.py)
    (33, 0, 0, 1, 0, 0.3, 12.3, 10_000, False, 0),        # nms = 0, threshold not representable in fp32
    (33, 0, 0, 3, 3, 0.3, 40.0, 10_000, False, 0),        # nms = 3 (49 cells), everything below 100 px is a candidate
    (33, 0, 0, 2, 2, 0.3, 16.0, 70, False, 0),            # max_factors hit by the neighbourhood edges alone
    (20, 4, 9, 2, 1, 0.3, 30.0, 500, True, 5),            # t1 > t0: (i,i) and (i,j) cells with j < t1 wrap like d[-k]
    (12, 12, 0, 2, 2, 0.3, 16.0, 100, False, 0),          # empty candidate rectangle
]


def problem(n_kf, seed, ht=24, wd=32):
    from droid_b200 import synth
    ii, jj = synth.neighborhood_edges(0, n_kf, 2)
    return synth.make_ba_problem(ii, jj, n_kf, ht=ht, wd=wd, seed=seed)


def case_inputs(case):
    n_kf, t0, t1, rad, nms, beta, thresh, max_factors, stereo, n_old = case
    p = problem(n_kf, 100 + n_kf + nms)
    rng = np.random.default_rng(5)
    io = rng.integers(0, n_kf, n_old).astype(np.int64); jo = rng.integers(0, n_kf, n_old).astype(np.int64)
    return p, io, jo


def filter_case():
    rng = np.random.default_rng(3)
    return rng.integers(0, 30, 500), rng.integers(0, 30, 500), rng.integers(0, 30, 300), rng.integers(0, 30, 300)
