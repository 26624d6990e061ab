"""Golden vectors produced by the REFERENCE's own CUDA extension on a B200 (tests/golden/make_ref_cuda_golden.py,
fixture tests/golden/ref_cuda_golden.npz).  CPU tests pin the oracle to them; gpu tests pin the sm_100a kernels."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
Z = np.load(os.path.join(HERE, "golden", "ref_cuda_golden.npz"))


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def bits(a):
    return a.view(np.uint16 if a.dtype == np.float16 else np.uint32)


def _fd_inputs():
    from droid_b200 import synth
    p = synth.config_c2()
    n = p.t1
    ii, jj = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    d = p.disps.copy(); d[3] = 1e3
    return p, d, ii.reshape(-1).astype(np.int64), jj.reshape(-1).astype(np.int64)


# ------------------------------------------------------------------------------------------ oracle (CPU)
@pytest.mark.parametrize("case", ["c1", "c1s", "c2", "c2b"])
def test_oracle_ba_vs_reference_cuda(case):
    import make_ref_cuda_golden as G
    from oracle import oracle
    p, lm, ep, _ = G.ba_cases()[case]
    # one Gauss-Newton iteration: 1e-4 normwise (north_star); two chained iterations: fp32 round-off compounds
    for it, tol in ((1, 1e-4), (2, 3e-4)):
        r = oracle.ba(p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, p.jj, p.t0, p.t1, it, lm, ep)
        assert rel(r["dx"], Z[f"ba_{case}_it{it}_dx"]) <= tol
        assert rel(r["dz"], Z[f"ba_{case}_it{it}_dz"]) <= tol
        assert np.abs(r["poses"][:p.t1] - Z[f"ba_{case}_it{it}_poses"]).max() <= 5e-6


def test_oracle_moba_vs_reference_cuda():
    import make_ref_cuda_golden as G
    from oracle import oracle
    p = G.moba_case()
    r = oracle.ba(p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, p.jj, p.t0, p.t1, 1, 1e-4, .1,
                  motion_only=True)
    assert rel(r["dx"], Z["moba_it1_dx"]) <= 1e-4
    assert np.abs(r["poses"] - Z["moba_it1_poses"]).max() <= 5e-6


@pytest.mark.parametrize("beta", [0.3, 0.25])
def test_oracle_frame_distance_bit_exact_vs_reference_cuda(beta):
    from oracle import oracle
    p, d, ii, jj = _fd_inputs()
    o = oracle.frame_distance(p.poses, d, p.intrinsics, ii, jj, beta)
    g = Z[f"fd_beta{beta}"]
    assert (g == 1000.0).any()
    np.testing.assert_array_equal(bits(o), bits(g))


@pytest.mark.parametrize("tag,shape,dt", [("f16", (2, 12, 16, 12, 16), np.float16), ("f32", (2, 12, 16, 12, 16), np.float32),
                                          ("f16b", (1, 24, 32, 6, 8), np.float16)])
def test_oracle_lookup_bit_exact_vs_reference_cuda(tag, shape, dt):
    import make_ref_cuda_golden as G
    from oracle import oracle
    vol, co = G.lookup_inputs(21, shape, dt)
    np.testing.assert_array_equal(bits(oracle.corr_index_forward(vol, co, 3)), bits(Z[f"lookup_{tag}"]))


def test_oracle_lookup_backward_bit_exact_vs_reference_cuda():
    import make_ref_cuda_golden as G
    from oracle import oracle
    vol, co = G.lookup_inputs(22, (2, 12, 16, 12, 16), np.float32)
    g = np.random.default_rng(23).normal(0, 1, (2, 7, 7, 12, 16)).astype(np.float32)
    np.testing.assert_array_equal(bits(oracle.corr_index_backward(vol.shape, co, g, 3)), bits(Z["lookup_bwd_f32"]))


def test_oracle_altcorr_vs_reference_cuda():
    import make_ref_cuda_golden as G
    from oracle import oracle
    f1, f2, co, g = G.altcorr_inputs(31)
    o = oracle.altcorr_forward(f1, f2, co, 3)
    assert np.abs(o - Z["altcorr_fwd"]).max() <= 1e-6 * np.abs(Z["altcorr_fwd"]).max()
    o1, o2, oc = oracle.altcorr_backward(f1, f2, co, g, 3)
    assert rel(o1, Z["altcorr_bwd_f1"]) <= 1e-6 and rel(o2, Z["altcorr_bwd_f2"]) <= 1e-6
    assert float(Z["altcorr_bwd_coords_absmax"]) == 0.0 and not oc.any()


# ------------------------------------------------------------------------------------------ kernels (GPU)
def _t(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["c1", "c1s", "c2", "c2b"])
def test_kernel_ba_vs_reference_cuda(case):
    import droid_backends
    import make_ref_cuda_golden as G
    p, lm, ep, _ = G.ba_cases()[case]
    for it, tol in ((1, 1e-4), (2, 3e-4)):
        poses, disps = _t(p.poses), _t(p.disps)
        dx, dz = droid_backends.ba(poses, disps, _t(p.intrinsics), _t(p.disps_sens), _t(p.target), _t(p.weight), _t(p.eta),
                                   _t(p.ii), _t(p.jj), p.t0, p.t1, it, lm, ep, False)
        assert rel(dx.cpu().numpy(), Z[f"ba_{case}_it{it}_dx"]) <= tol
        assert rel(dz.cpu().numpy(), Z[f"ba_{case}_it{it}_dz"]) <= tol
        assert np.abs(poses.cpu().numpy()[:p.t1] - Z[f"ba_{case}_it{it}_poses"]).max() <= 5e-6


@pytest.mark.gpu
def test_kernel_moba_vs_reference_cuda():
    import droid_backends
    import make_ref_cuda_golden as G
    p = G.moba_case()
    poses, disps = _t(p.poses), _t(p.disps)
    dx, dz = droid_backends.ba(poses, disps, _t(p.intrinsics), _t(p.disps_sens), _t(p.target), _t(p.weight), _t(p.eta),
                               _t(p.ii), _t(p.jj), p.t0, p.t1, 1, 1e-4, .1, True)
    assert dz is None and rel(dx.cpu().numpy(), Z["moba_it1_dx"]) <= 1e-4


@pytest.mark.gpu
@pytest.mark.parametrize("beta", [0.3, 0.25])
def test_kernel_frame_distance_bit_exact_vs_reference_cuda(beta):
    import droid_backends
    p, d, ii, jj = _fd_inputs()
    o = droid_backends.frame_distance(_t(p.poses), _t(d), _t(p.intrinsics), _t(ii), _t(jj), beta).cpu().numpy()
    np.testing.assert_array_equal(bits(o), bits(Z[f"fd_beta{beta}"]))


@pytest.mark.gpu
@pytest.mark.parametrize("tag,shape,dt", [("f16", (2, 12, 16, 12, 16), np.float16), ("f32", (2, 12, 16, 12, 16), np.float32),
                                          ("f16b", (1, 24, 32, 6, 8), np.float16)])
def test_kernel_lookup_bit_exact_vs_reference_cuda(tag, shape, dt):
    import droid_backends
    import make_ref_cuda_golden as G
    vol, co = G.lookup_inputs(21, shape, dt)
    o, = droid_backends.corr_index_forward(_t(vol), _t(co), 3)
    np.testing.assert_array_equal(bits(o.cpu().numpy()), bits(Z[f"lookup_{tag}"]))


@pytest.mark.gpu
def test_kernel_lookup_backward_and_altcorr_vs_reference_cuda():
    import droid_backends
    import make_ref_cuda_golden as G
    vol, co = G.lookup_inputs(22, (2, 12, 16, 12, 16), np.float32)
    g = np.random.default_rng(23).normal(0, 1, (2, 7, 7, 12, 16)).astype(np.float32)
    o, = droid_backends.corr_index_backward(_t(vol), _t(co), _t(g), 3)
    np.testing.assert_array_equal(bits(o.cpu().numpy()), bits(Z["lookup_bwd_f32"]))
    f1, f2, co, g = G.altcorr_inputs(31)
    o, = droid_backends.altcorr_forward(_t(f1), _t(f2), _t(co), 3)
    assert np.abs(o.cpu().numpy() - Z["altcorr_fwd"]).max() <= 1e-4 * np.abs(Z["altcorr_fwd"]).max()
    g1, g2, gc = droid_backends.altcorr_backward(_t(f1), _t(f2), _t(co), _t(g), 3)
    assert rel(g1.cpu().numpy(), Z["altcorr_bwd_f1"]) <= 1e-4 and rel(g2.cpu().numpy(), Z["altcorr_bwd_f2"]) <= 1e-4


# ---------------------------------------------------------------- second fixture: projmap, iproj, depth_filter
Z2 = np.load(os.path.join(HERE, "golden", "ref_cuda_golden2.npz"))


def _extra():
    import make_ref_cuda_golden as G
    p = G.extra_problem()
    sel = p.ii != p.jj
    return p, p.ii[sel], p.jj[sel]


def test_oracle_projmap_iproj_depth_filter_vs_reference_cuda():
    from oracle import oracle
    p, ii, jj = _extra()
    c, v = oracle.projmap(p.poses, p.disps, p.intrinsics, ii, jj)
    np.testing.assert_allclose(c, Z2["projmap_coords"], rtol=1e-4, atol=1e-2)      # coordinates reach ~1e3 px near Z = 0.01
    np.testing.assert_array_equal(v, Z2["projmap_valid"])
    np.testing.assert_allclose(oracle.iproj(p.poses, p.disps, p.intrinsics), Z2["iproj"], rtol=1e-4, atol=1e-5)
    n = p.poses.shape[0]
    df = oracle.depth_filter(p.poses, p.disps, p.intrinsics, np.arange(n), np.full(n, 0.05, np.float32))
    np.testing.assert_array_equal(df, Z2["depth_filter"])                           # integer counts


@pytest.mark.gpu
def test_kernel_projmap_iproj_depth_filter_vs_reference_cuda():
    import droid_backends
    p, ii, jj = _extra()
    c, v = droid_backends.projmap(_t(p.poses), _t(p.disps), _t(p.intrinsics), _t(ii), _t(jj))
    np.testing.assert_allclose(c.cpu().numpy(), Z2["projmap_coords"], rtol=1e-4, atol=1e-2)
    np.testing.assert_array_equal(v.cpu().numpy(), Z2["projmap_valid"])
    pts = droid_backends.iproj(_t(p.poses), _t(p.disps), _t(p.intrinsics))
    np.testing.assert_allclose(pts.cpu().numpy(), Z2["iproj"], rtol=1e-4, atol=1e-5)
    n = p.poses.shape[0]
    import torch
    df = droid_backends.depth_filter(_t(p.poses), _t(p.disps), _t(p.intrinsics), torch.arange(n).cuda(),
                                     torch.full((n,), 0.05).cuda())
    np.testing.assert_array_equal(df.cpu().numpy(), Z2["depth_filter"])


# ---------------------------------------------------------------- third fixture: the configs the metric is quoted on
# tests/golden/make_c3_golden.py: reference CUDA extension (ref_cuda_c3.npz, B200) and oracle fp32/fp64 (oracle_c3.npz, CPU)
# on c3 (512 KF / ~6000 edges / 48x64), c3s (20 % stereo edges + sensor prior, 40x64) and c4 (1536 KF / ~20k edges).
# What "within tolerance of the reference" can mean at this size is MEASURED, not argued: the reference's own fp32 result
# sits 5e-4 (c3) .. 1.4e-3 (c3s, 2 iterations) away from exact arithmetic (the fp64 oracle) in dx, so no implementation
# with a different summation order can be within 1e-4 of it.  Asserted instead, per call:
#   (a) the kernel is at least as close to exact arithmetic as the reference is: d(kernel, f64) <= max(1e-4, d(ref, f64));
#   (b) d(kernel, ref) <= max(1e-4, 2 d(ref, f64))   (triangle inequality with (a));
#   (c) dz (well conditioned) within 1e-4 of the reference for one iteration, d(kernel, f64)-bounded after two.
ZR = np.load(os.path.join(HERE, "golden", "ref_cuda_c3.npz"))
ZO = np.load(os.path.join(HERE, "golden", "oracle_c3.npz"))


def _c3_cases():
    import make_c3_golden as G3
    return G3


def test_oracle_c3_vs_reference_cuda_three_way_on_record():
    """CPU: reference-CUDA vs oracle distances at c3/c3s — the reference's own fp32 noise level at the headline size."""
    for name in ("c3", "c3s"):
        for it in (1, 2):
            ref, o32, o64 = ZR[f"{name}_it{it}_ref_dx"], ZO[f"{name}_it{it}_f32_dx"], ZO[f"{name}_it{it}_f64_dx"]
            d_ref64, d_3264 = rel(ref, o64), rel(o32, o64)
            assert 1e-5 < d_ref64 < 5e-3 and d_3264 < 5e-3, (name, it, d_ref64, d_3264)
            assert rel(ref, o32) <= 2 * max(d_ref64, d_3264)
            assert rel(ZR[f"{name}_it{it}_ref_dz_sample"], ZO[f"{name}_it{it}_f64_dz_sample"]) <= 3e-4


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["c3", "c3s", "c4"])
def test_kernel_ba_c3_c4_vs_reference_cuda(name):
    import droid_backends
    G3 = _c3_cases()
    p, its = G3.cases((name,))[name]
    args = [_t(p.intrinsics), _t(p.disps_sens), _t(p.target), _t(p.weight), _t(p.eta), _t(p.ii), _t(p.jj)]
    for it in its:
        poses, disps = _t(p.poses), _t(p.disps)
        dx, dz = droid_backends.ba(poses, disps, *args, p.t0, p.t1, it, G3.LM, G3.EP, False)
        dx, dz = dx.cpu().numpy(), dz.cpu().numpy()
        ref_dx, o64_dx = ZR[f"{name}_it{it}_ref_dx"], ZO[f"{name}_it{it}_f64_dx"]
        r, c = G3.dz_sample_index(*dz.shape)
        d_ref64 = rel(ref_dx, o64_dx)
        d_k64, d_kref = rel(dx, o64_dx), rel(dx, ref_dx)
        dz_kref = rel(dz[r, c], ZR[f"{name}_it{it}_ref_dz_sample"])
        dz_k64 = rel(dz[r, c], ZO[f"{name}_it{it}_f64_dz_sample"])
        dz_ref64 = rel(ZR[f"{name}_it{it}_ref_dz_sample"], ZO[f"{name}_it{it}_f64_dz_sample"])
        print(f"{name} it{it}: dx kernel-f64 {d_k64:.2e} ref-f64 {d_ref64:.2e} kernel-ref {d_kref:.2e}; "
              f"dz kernel-ref {dz_kref:.2e} kernel-f64 {dz_k64:.2e} ref-f64 {dz_ref64:.2e}")
        assert d_k64 <= max(1e-4, d_ref64), (d_k64, d_ref64)
        assert d_kref <= max(1e-4, 2 * d_ref64), (d_kref, d_ref64)
        assert dz_k64 <= max(1e-4, 2 * dz_ref64) and dz_kref <= max(1e-4, 2 * max(dz_ref64, dz_k64))
        assert abs(np.linalg.norm(dz.astype(np.float64)) / float(ZR[f"{name}_it{it}_ref_dz_norm"]) - 1) <= 1e-3
        rn = np.linalg.norm(dz.astype(np.float64), axis=1)
        assert np.abs(rn - ZR[f"{name}_it{it}_ref_dz_rownorm"]).max() <= 2e-3 * ZR[f"{name}_it{it}_ref_dz_rownorm"].max()
