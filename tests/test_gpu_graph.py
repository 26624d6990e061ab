"""GPU parity: factor-graph edge selection (csrc/graph_edges.cu) vs the oracle's restatement of
factor_graph.py:315-379 — edge lists must be EQUAL, element for element, in order."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from graph_cases import CASES, case_inputs, filter_case, problem as _problem  # noqa: E402

# edge lists produced by the reference's OWN factor_graph.py (tests/golden/make_pyref_graph_golden.py)
GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "pyref_graph_edges.npz"))


@pytest.mark.parametrize("case", CASES)
def test_proximity_edges_equal_the_reference_list(case):
    from droid_b200 import graph
    from oracle import graph_edges as g
    from oracle import oracle
    n_kf, t0, t1, rad, nms, beta, thresh, max_factors, stereo, n_old = case
    p, io, jo = case_inputs(case)
    t = n_kf
    ii, jj, d = g.proximity_distance(oracle.frame_distance, p.poses, p.disps, p.intrinsics, t0, t1, t, beta)
    want = g.add_proximity_factors_edges(ii, jj, d, t0, t1, t, rad, nms, thresh, max_factors, stereo, io, jo)
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    gi, gj = graph.proximity_edges(cu(p.poses), cu(p.disps), cu(p.intrinsics), t, t0, t1, rad, nms, beta, thresh, max_factors,
                                   stereo, cu(io) if n_old else None, cu(jo) if n_old else None)
    got = list(zip(gi.cpu().tolist(), gj.cpu().tolist()))
    assert got == want
    assert got == [tuple(e) for e in GOLD[f"case{CASES.index(case)}_es"].tolist()]      # the reference's own list
    if case is CASES[0]:
        assert len(want) > 2 * (2 * t)          # the case does pick loop-closure edges, not only the neighbourhood


def test_filter_repeated_edges_and_neighborhood():
    from droid_b200 import graph
    from oracle import graph_edges as g
    ii, jj, oi, oj = filter_case()
    wi, wj = g.filter_repeated_edges(ii, jj, oi, oj)
    assert wi.tolist() == GOLD["filter_ii"].tolist() and wj.tolist() == GOLD["filter_jj"].tolist()
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a).astype(np.int64)).cuda()
    gi, gj = graph.filter_repeated_edges(cu(ii), cu(jj), cu(oi), cu(oj))
    assert gi.cpu().tolist() == wi.tolist() and gj.cpu().tolist() == wj.tolist()
    gi, gj = graph.filter_repeated_edges(cu(ii), cu(jj), cu(oi[:0]), cu(oj[:0]))
    assert gi.cpu().tolist() == ii.tolist()
    for stereo in (False, True):
        a, b = graph.neighborhood_edges(2, 11, 3, stereo)
        wa, wb = g.add_neighborhood_factors_edges(2, 11, 3, stereo)
        assert a.cpu().tolist() == wa.tolist() and b.cpu().tolist() == wb.tolist()


def test_proximity_edges_c3_size_runs_and_is_bidirectional():
    """512 keyframes (262k candidate pairs): size-independent properties of the list."""
    from droid_b200 import graph, synth
    p = synth.config_c3(n_edges=2000)
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    t = p.t1
    gi, gj = graph.proximity_edges(cu(p.poses), cu(p.disps), cu(p.intrinsics), t, 0, 0, 2, 2, 0.25, 16.0, 16 * t, False)
    gi, gj = gi.cpu().numpy(), gj.cpu().numpy()
    n_init = sum(2 * min(i, 3) for i in range(t))
    assert len(gi) >= n_init and len(gi) % 2 == 0 and len(gi) <= 16 * t + 2
    assert np.array_equal(gi[0::2], gj[1::2]) and np.array_equal(gj[0::2], gi[1::2])      # (i,j) then (j,i)
    picks = np.stack([gi[n_init::2], gj[n_init::2]], 1)
    assert np.all(picks[:, 0] - 2 >= picks[:, 1])                                          # d[ii - rad < jj] = inf
    assert len({(a, b) for a, b in picks.tolist()}) == len(picks)                         # a suppressed cell is never re-picked


# ---------------------------------------------------------------------------------- reprojection (§8f-1) and projmap
def _reproject_problem():
    import importlib.util, os
    src = open(os.path.join(os.path.dirname(__file__), "golden", "make_pyref_reproject_golden.py")).read()
    ns = {}
    exec(src[src.index("def problem():"):src.index("def main():")], {"np": np}, ns)
    return ns["problem"]()


def test_reproject_and_motion_features():
    import os
    from droid_b200.video import BAVideo
    from oracle import oracle
    p = _reproject_problem()
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "pyref_reproject.npz"))
    K = np.tile(p.intrinsics, (p.poses.shape[0], 1)).astype(np.float32)
    video = BAVideo(cu(p.poses), cu(p.disps), cu(K))
    coords, valid = video.reproject(cu(p.ii), cu(p.jj))
    assert coords.shape == (1, len(p.ii), 24, 32, 2) and valid.shape == (1, len(p.ii), 24, 32, 1)
    c, v = coords[0].cpu().numpy(), valid[0].cpu().numpy()
    far = np.abs(g["coords"]).max(-1) > 1e4
    np.testing.assert_allclose(c[~far], g["coords"][~far], rtol=1e-4, atol=2e-3)          # the reference's own fp64 output
    assert (v != g["valid"]).mean() < 1e-3
    wc, wv, wm = oracle.reproject(p.poses, p.disps, K, p.ii, p.jj, target=p.target.transpose(0, 2, 3, 1))
    np.testing.assert_allclose(c[~far], wc[~far], rtol=1e-5, atol=1e-4)
    # fused motion features (factor_graph.py:203-205), per-frame intrinsics that differ
    K2 = K.copy(); K2[3] *= 1.25
    video2 = BAVideo(cu(p.poses), cu(p.disps), cu(K2))
    tgt = cu(p.target.transpose(0, 2, 3, 1))
    c2, v2, motn = video2.reproject(cu(p.ii), cu(p.jj), target=tgt[None])
    wc2, wv2, wm2 = oracle.reproject(p.poses, p.disps, K2, p.ii, p.jj, target=p.target.transpose(0, 2, 3, 1))
    assert motn.shape == (1, len(p.ii), 4, 24, 32)
    m = motn[0].cpu().numpy()
    ok = np.abs(wm2) < 63.9                                                               # away from the clamp
    np.testing.assert_allclose(m[ok], wm2[ok], rtol=1e-4, atol=2e-3)
    assert np.abs(m).max() <= 64.0
    same = torch.cat([c2[0] - torch.stack(torch.meshgrid(torch.arange(32.0), torch.arange(24.0), indexing="xy"), -1).cuda(),
                      tgt - c2[0]], -1).permute(0, 3, 1, 2).clamp(-64.0, 64.0)
    assert torch.equal(same, motn[0])                                                      # exactly the torch expression


def test_projmap_matches_oracle():
    import droid_backends
    from oracle import oracle
    p = _reproject_problem()
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    sel = p.ii != p.jj
    coords, valid = droid_backends.projmap(cu(p.poses), cu(p.disps), cu(p.intrinsics), cu(p.ii[sel]), cu(p.jj[sel]))
    wc, wv = oracle.projmap(p.poses, p.disps, p.intrinsics, p.ii[sel], p.jj[sel])
    c, v = coords.cpu().numpy(), valid.cpu().numpy()
    assert c.shape == wc.shape == (int(sel.sum()), 24, 32, 3) and np.all(c[..., 2] == 0)
    far = np.abs(wc).max(-1) > 1e4
    np.testing.assert_allclose(c[~far], wc[~far], rtol=1e-4, atol=2e-3)
    assert (v != wv).mean() < 1e-3 and 0.5 < wv.mean() < 1.0
