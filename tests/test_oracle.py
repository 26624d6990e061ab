"""CPU tests of the oracle itself: pinned to the reference's PyTorch BA (fp64 golden vectors), Eigen LLT
cross-check, structural properties of the lookups, and the C-ABI library's exported symbols."""
import os
import re
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(HERE, "golden"))


@pytest.mark.parametrize("name", ["nbr6", "back10"])
def test_oracle_matches_reference_pytorch_ba(name):
    """oracle (fp64, CUDA/PyTorch divergences switched off) == geom/ba.py BA and MoBA to 1e-9."""
    import make_pyref_golden as g
    from oracle import oracle
    p = g.cases()[name]
    z = np.load(os.path.join(HERE, "golden", f"pyref_ba_{name}.npz"))
    poses = p.poses.astype(np.float64)
    poses[:, 3:] /= np.linalg.norm(poses[:, 3:], axis=1, keepdims=True)
    t1 = p.t1
    for it in (1, 2):
        r = oracle.ba(poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta.astype(np.float64) + 1e-7, p.ii,
                      p.jj, p.t0, p.t1, it, 1e-4, 0.1, skip_pose0_backsub=False, damp_before_schur=True, dtype=np.float64)
        assert np.abs(r["poses"][:t1] - z[f"poses_it{it}"]).max() < 1e-9
        assert np.abs(r["disps"][:t1] - z[f"disps_it{it}"]).max() < 1e-9
        assert np.abs(z[f"poses_it{it}"] - poses[:t1]).max() > 1e-3     # the fixture really moved
    r = oracle.ba(poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, None, p.ii, p.jj, p.t0, p.t1, 2, 1e-4, 0.1,
                  motion_only=True, dtype=np.float64, moba_block_damping=True)
    assert np.abs(r["poses"][:t1] - z["poses_moba2"]).max() < 1e-9


def test_dense_cholesky_equals_eigen_simplicial_llt():
    """the oracle's dense fp64 solve vs SparseBlock::solve through the reference's vendored Eigen (oracle/_ref)."""
    from droid_b200 import synth
    from oracle import oracle
    p = synth.config_c1()
    r = oracle.ba(p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, p.jj, p.t0, p.t1, 1, 1e-4, 0.1,
                  dump_system=True)
    res = oracle.ref_eigen_solve(r["H"], r["b"], 1e-4, 0.1)
    if res is None:
        pytest.skip("oracle/_ref not built (needs /root/reference/thirdparty/eigen)")
    x, st = res
    assert st == 0
    np.testing.assert_allclose(x.astype(np.float32).reshape(-1, 6), r["dx"], rtol=1e-6, atol=1e-9)
    x, st = oracle.ref_eigen_solve(-r["H"], r["b"], 0.0, 0.0)
    assert st == 1 and not x.any()      # zeros on failure (droid_kernels.cu:1207-1210)


def test_ba_oracle_fp32_vs_fp64_spread_and_convergence():
    from droid_b200 import synth
    from oracle import oracle
    p = synth.config_c1()
    a = oracle.ba(p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, p.jj, p.t0, p.t1, 1, 1e-4, 0.1)
    b = oracle.ba(p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, p.jj, p.t0, p.t1, 1, 1e-4, 0.1,
                  dtype=np.float64)
    assert np.linalg.norm(a["dx"] - b["dx"]) / np.linalg.norm(b["dx"]) < 1e-4
    r = oracle.ba(p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, p.jj, p.t0, p.t1, 6, 1e-4, 0.1)
    gt = synth.camera_centres(p.poses_gt[:8])
    assert synth.umeyama_ate(synth.camera_centres(r["poses"][:8]), gt) < 0.2 * synth.umeyama_ate(synth.camera_centres(p.poses[:8]), gt)
    with pytest.raises(RuntimeError):
        jj = p.jj.copy(); jj[0] = 50
        oracle.ba(p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, jj, p.t0, p.t1, 1, 1e-4, 0.1)


def test_jacobians_by_finite_differences():
    """dx from one fp64 Gauss-Newton step must reduce the weighted reprojection cost."""
    from droid_b200 import synth
    from oracle import oracle
    p = synth.config_c1()

    def cost(poses, disps):
        c, z = synth.reproject(poses.astype(np.float64), disps.astype(np.float64), p.intrinsics, p.ii, p.jj)
        r = p.target.transpose(0, 2, 3, 1) - c
        return float((0.001 * p.weight.transpose(0, 2, 3, 1) * r ** 2).sum())

    r = oracle.ba(p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, p.jj, p.t0, p.t1, 1, 1e-4, 0.1,
                  dtype=np.float64)
    assert cost(r["poses"], r["disps"]) < 0.2 * cost(p.poses, p.disps)


def test_lookup_at_integer_coords_is_window_copy():
    from oracle import oracle
    rng = np.random.default_rng(0)
    vol = rng.normal(0, 1, (1, 4, 5, 10, 12)).astype(np.float32)
    co = np.zeros((1, 2, 4, 5), np.float32); co[0, 0] = 6.0; co[0, 1] = 4.0
    out = oracle.corr_index_forward(vol, co, 3)
    for i in range(7):
        for j in range(7):       # first output axis is the x offset (correlation_kernels.cu:49-66)
            np.testing.assert_array_equal(out[0, i, j], vol[0, :, :, 4 - 3 + j, 6 - 3 + i])
    co[:] = -100.0
    assert not oracle.corr_index_forward(vol, co, 3).any()                     # fully outside -> zeros
    assert oracle.corr_index_forward(vol[:0], co[:0], 3).shape == (0, 7, 7, 4, 5)   # empty batch


def test_lookup_backward_is_transpose_of_forward():
    from oracle import oracle
    rng = np.random.default_rng(1)
    shape = (2, 5, 6, 7, 9)
    vol = rng.normal(0, 1, shape).astype(np.float32)
    co = np.stack([rng.uniform(-2, 11, (2, 5, 6)), rng.uniform(-2, 9, (2, 5, 6))], 1).astype(np.float32)
    g = rng.normal(0, 1, (2, 7, 7, 5, 6)).astype(np.float32)
    lhs = float((oracle.corr_index_forward(vol, co, 3).astype(np.float64) * g).sum())
    rhs = float((oracle.corr_index_backward(shape, co, g, 3).astype(np.float64) * vol).sum())
    assert abs(lhs - rhs) <= 1e-4 * abs(lhs)


def test_altcorr_equals_lookup_of_volume_and_adjoint():
    from oracle import oracle
    rng = np.random.default_rng(2)
    N, C, h, w = 2, 32, 6, 8
    f1 = rng.normal(0, 1, (N, C, h, w)).astype(np.float32)
    f2 = rng.normal(0, 1, (N, C, h, w)).astype(np.float32)
    co = np.stack([rng.uniform(-2, w + 2, (N, h, w)), rng.uniform(-2, h + 2, (N, h, w))], 1).astype(np.float32)
    vol = oracle.corr_volume(f1, f2, 1)[0]
    a = oracle.corr_index_forward(vol, co, 3).reshape(N, 49, h, w)
    cl = lambda f: np.ascontiguousarray((f / 4).transpose(0, 2, 3, 1))
    co5 = np.ascontiguousarray(co.transpose(0, 2, 3, 1))[:, None]
    b = oracle.altcorr_forward(cl(f1), cl(f2), co5, 3)[:, 0]
    np.testing.assert_allclose(a, b, rtol=0, atol=1e-5)
    g = rng.normal(0, 1, (N, 1, 49, h, w)).astype(np.float32)
    g1, g2, gc = oracle.altcorr_backward(cl(f1), cl(f2), co5, g, 3)
    lhs = float((b[:, None].astype(np.float64) * g).sum())
    assert abs(lhs - float((g1.astype(np.float64) * cl(f1)).sum())) <= 1e-4 * abs(lhs)   # bilinear in (f1,f2)
    assert abs(lhs - float((g2.astype(np.float64) * cl(f2)).sum())) <= 1e-4 * abs(lhs)
    assert not gc.any()


def test_corr_pyramid_levels_are_average_pools():
    from oracle import oracle
    rng = np.random.default_rng(3)
    f1 = rng.normal(0, 1, (1, 16, 4, 8)).astype(np.float32)
    f2 = rng.normal(0, 1, (1, 16, 4, 8)).astype(np.float32)
    lv = oracle.corr_volume(f1, f2, 3)
    ref0 = np.einsum("ncp,ncq->npq", f1.reshape(1, 16, 32) / 4, f2.reshape(1, 16, 32) / 4).reshape(1, 4, 8, 4, 8)
    np.testing.assert_allclose(lv[0], ref0, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(lv[1], lv[0].reshape(1, 4, 8, 2, 2, 4, 2).mean((4, 6)), rtol=1e-5, atol=1e-6)
    assert lv[2].shape == (1, 4, 8, 1, 2)


def test_se3_properties():
    """lietorch's group property tests restated for the oracle's SE3 (run_tests.py:16-54): Exp(0)=I, retraction by
    xi then -xi returns, relative pose of a pose with itself is identity."""
    from oracle import oracle
    rng = np.random.default_rng(4)
    pose = np.concatenate([rng.normal(0, 1, 3), [0.1, -0.2, 0.3, 0.0]]).astype(np.float32)
    pose[3:] /= np.linalg.norm(pose[3:])
    np.testing.assert_allclose(oracle.se3_retr(np.zeros(6, np.float32), pose), pose, atol=1e-7)
    xi = rng.normal(0, 0.1, 6).astype(np.float32)
    back = oracle.se3_retr(-xi, oracle.se3_retr(xi, pose))
    np.testing.assert_allclose(back, pose, atol=2e-6)
    np.testing.assert_allclose(oracle.se3_rel(pose, pose), [0, 0, 0, 0, 0, 0, 1], atol=1e-6)


def test_frame_distance_identity_and_invalid():
    from droid_b200 import synth
    from oracle import oracle
    p = synth.config_c1()
    ii = np.arange(8, dtype=np.int64)
    d = oracle.frame_distance(p.poses, p.disps, p.intrinsics, ii, ii, 0.3)
    assert np.all(d < 1e-4)                                   # a frame is at distance ~0 from itself
    tz = oracle.se3_rel(p.poses[0], p.poses[1])[2]
    disps = p.disps.copy(); disps[0] = -1e3 * np.sign(tz)     # every point of frame 0 lands behind camera 1
    assert oracle.frame_distance(p.poses, disps, p.intrinsics, np.array([0]), np.array([1]), 0.3)[0] == 1000.0
    assert oracle.frame_distance(p.poses, p.disps, p.intrinsics, ii[:0], ii[:0], 0.3).shape == (0,)


# ------------------------------------------------------------------------------------------ C-ABI library
def test_capi_library_exports_every_declared_symbol():
    from droid_b200 import _capi
    hdr = open(os.path.join(REPO, "include", "droid_b200.h")).read()
    declared = set(re.findall(r"\b(db200_\w+)\s*\(", hdr))
    assert declared, "no declarations found"
    lib = _capi.lib()
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/droid_b200.h but not exported"
    assert set(_capi.EXPORTED) <= declared
    assert lib.db200_version() >= 100
    assert lib.db200_ba_workspace_bytes(36, 8, 48, 64, 1, 8, 8, 0) > 0
    assert lib.db200_ba_workspace_bytes(36, 8, 48, 64, 3, 3, 8, 0) == 0
    assert b"eta" in lib.db200_status_string(-5)


def test_droid_backends_surface_and_no_cpu_fallback():
    import torch
    import droid_backends
    for name in ["ba", "frame_distance", "projmap", "depth_filter", "iproj", "altcorr_forward", "altcorr_backward",
                 "corr_index_forward", "corr_index_backward"]:            # droid.cpp:237-250
        assert callable(getattr(droid_backends, name))
    with pytest.raises(RuntimeError):                                     # CPU tensors are rejected, never emulated
        droid_backends.corr_index_forward(torch.zeros(1, 2, 2, 2, 2), torch.zeros(1, 2, 2, 2), 3)
    with pytest.raises(RuntimeError):                                     # CHECK_CONTIGUOUS (droid.cpp:84-85)
        droid_backends.corr_index_forward(torch.zeros(1, 2, 2, 2, 4)[..., ::2], torch.zeros(1, 2, 2, 2), 3)


def test_tiled_slice_layout_helpers_match_the_header_formula():
    """include/droid_b200.h, "tiled levels": element (y, x) of a [h2][w2] slice sits at ((y>>2)*(w2>>3) + (x>>3))*32 + (y&3)*8 + (x&7).
    `tile_level` / `untile_level` are the host-side converters (pure tensor reshapes: no GPU needed) and are inverse to each other."""
    import torch
    import droid_backends
    for h2, w2 in [(48, 64), (24, 32), (8, 40)]:
        v = torch.arange(3 * h2 * w2, dtype=torch.int32).reshape(3, 1, h2, w2)
        t = droid_backends.tile_level(v)
        assert t.shape == v.shape and torch.equal(droid_backends.untile_level(t), v)
        flat = t.reshape(3, -1)
        for y, x in [(0, 0), (3, 7), (4, 8), (h2 - 1, w2 - 1), (5, 17), (h2 // 2, w2 // 2 + 3)]:
            idx = ((y >> 2) * (w2 >> 3) + (x >> 3)) * 32 + (y & 3) * 8 + (x & 7)
            assert int(flat[1, idx]) == int(v[1, 0, y, x]), (h2, w2, y, x)
    assert droid_backends.corr_tiled_levels(128, 48, 64, 4, torch.float64) == 0      # fp16 only


# ------------------------------------------------------------------------------------------ edge selection (§8f-2)
def test_proximity_edges_oracle_hand_worked_case():
    """6 frames, rad=1, nms=1: 18 neighbourhood edges, then picks (5,0), (4,1), (3,0), (5,2) — (4,0), (5,1) fall to the
    suppression of earlier picks (factor_graph.py:355-378 worked by hand)."""
    from oracle import graph_edges as g
    t = 6
    ii, jj = np.meshgrid(np.arange(t), np.arange(t), indexing="ij")
    ii, jj = ii.reshape(-1), jj.reshape(-1)
    d = np.full(t * t, 5.0, np.float32); d[5 * 6 + 0] = 1.0; d[4 * 6 + 1] = 2.0
    es = g.add_proximity_factors_edges(ii, jj, d, 0, 0, t, 1, 1, 16.0, 100, False, [], [])
    nb = [(1, 0), (0, 1), (2, 0), (0, 2), (2, 1), (1, 2), (3, 1), (1, 3), (3, 2), (2, 3), (4, 2), (2, 4), (4, 3), (3, 4),
          (5, 3), (3, 5), (5, 4), (4, 5)]
    assert es == nb + [(5, 0), (0, 5), (4, 1), (1, 4), (3, 0), (0, 3), (5, 2), (2, 5)]
    # max_factors: the check precedes each pick (len(es) > max_factors): 18 <= 19 lets one pick in, 20 > 19 stops
    es = g.add_proximity_factors_edges(ii, jj, d, 0, 0, t, 1, 1, 16.0, 19, False, [], [])
    assert es == nb + [(5, 0), (0, 5)]
    es = g.add_proximity_factors_edges(ii, jj, d, 0, 0, t, 1, 1, 16.0, 17, False, [], [])
    assert es == nb
    # threshold and existing edges: (5,0) too far; an existing (4,1) suppresses itself, (3,1), (5,1), (4,0), (4,2)
    d2 = d.copy(); d2[5 * 6 + 0] = 17.0
    es = g.add_proximity_factors_edges(ii, jj, d2, 0, 0, t, 1, 1, 16.0, 100, False, [4], [1])
    assert es == nb + [(3, 0), (0, 3), (5, 2), (2, 5)]
    # stereo: (i,i) first for every frame
    es = g.add_proximity_factors_edges(ii, jj, d, 0, 0, t, 1, 1, 0.5, 100, True, [], [])
    assert es[:3] == [(0, 0), (1, 1), (1, 0)] and len(es) == 18 + 6


def test_filter_repeated_and_neighborhood_edges_oracle():
    from oracle import graph_edges as g
    ii, jj = g.add_neighborhood_factors_edges(0, 8, 3)
    assert len(ii) == 36 and np.all(np.abs(ii - jj) <= 3) and np.all(ii != jj)        # BASELINE config 1: 36 edges
    ii, jj = g.add_neighborhood_factors_edges(0, 8, 3, stereo=True)
    assert len(ii) == 22                                                               # stereo variant
    a, b = g.filter_repeated_edges([1, 2, 2, 3], [0, 1, 1, 5], [2, 9], [1, 9])
    assert a.tolist() == [1, 3] and b.tolist() == [0, 5]


def test_reproject_oracle_matches_reference_projective_transform():
    """oracle.reproject (fp32) vs the reference's own projective_ops.projective_transform run in fp64 under the
    lietorch shim (tests/golden/make_pyref_reproject_golden.py): stereo edges, near points, a non-unit quaternion."""
    import importlib.util
    from oracle import oracle
    spec = importlib.util.spec_from_file_location("mk", os.path.join(REPO, "tests", "golden", "make_pyref_reproject_golden.py"))
    g = np.load(os.path.join(REPO, "tests", "golden", "pyref_reproject.npz"))
    src = open(spec.origin).read()
    ns = {}
    exec(src[src.index("def problem():"):src.index("def main():")], {"np": np}, ns)      # the seeded problem only
    p = ns["problem"]()
    coords, valid = oracle.reproject(p.poses, p.disps, p.intrinsics, p.ii, p.jj)
    assert 0.5 < g["valid"].mean() < 1.0                                                  # the case has invalid points
    far = np.abs(g["coords"]).max(-1) > 1e4                                               # |Z| ~ 0: ill-conditioned pixels
    np.testing.assert_allclose(coords[~far], g["coords"][~far], rtol=1e-4, atol=2e-3)
    assert (valid != g["valid"]).mean() < 1e-3                                            # depth exactly at the 0.2 threshold
    K = np.tile(p.intrinsics, (p.poses.shape[0], 1)); K[3] *= 1.25
    c2, _, motn = oracle.reproject(p.poses, p.disps, K, p.ii, p.jj, target=g["coords"])
    sel = p.ii != 3
    np.testing.assert_array_equal(c2[sel & (p.jj != 3)], coords[sel & (p.jj != 3)])
    assert motn.shape == (len(p.ii), 4, 24, 32) and np.abs(motn).max() <= 64.0


def test_edge_selection_oracle_equals_the_reference_factor_graph_lists():
    """oracle/graph_edges.py against the lists produced by the reference's own `add_proximity_factors`,
    `__filter_repeated_edges` and `add_neighborhood_factors` (factor_graph.py:44-55,302-379, run unmodified by
    tests/golden/make_pyref_graph_golden.py): equal element for element on all 8 cases."""
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import graph_cases as GC
    from oracle import graph_edges as g
    from oracle import oracle
    gold = np.load(os.path.join(REPO, "tests", "golden", "pyref_graph_edges.npz"))
    for c, case in enumerate(GC.CASES):
        n_kf, t0, t1, rad, nms, beta, thresh, max_factors, stereo, n_old = case
        p, io, jo = GC.case_inputs(case)
        ii, jj, d = g.proximity_distance(oracle.frame_distance, p.poses, p.disps, p.intrinsics, t0, t1, n_kf, beta)
        es = g.add_proximity_factors_edges(ii, jj, d, t0, t1, n_kf, rad, nms, thresh, max_factors, stereo, io, jo)
        assert es == [tuple(e) for e in gold[f"case{c}_es"].tolist()], c
    assert len(gold["case0_es"]) > 4 * 40 and str(gold["case7_reference_raised"]) == "ValueError"   # empty list: the reference raises
    ii, jj, oi, oj = GC.filter_case()
    wi, wj = g.filter_repeated_edges(ii, jj, oi, oj)
    assert wi.tolist() == gold["filter_ii"].tolist() and wj.tolist() == gold["filter_jj"].tolist()
    for stereo in (False, True):
        ni, nj = g.add_neighborhood_factors_edges(2, 11, 3, stereo)
        assert np.stack([ni, nj], 1).tolist() == gold[f"nbr_stereo{int(stereo)}"].tolist()
