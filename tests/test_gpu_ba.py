"""GPU parity: dense BA and frame_distance vs the CPU oracle (through droid_backends -> C-ABI)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

# north_star: per-iteration pose/disparity updates within 1e-4 relative error of the reference.
# Measured normwise: ||x - x_ref|| / ||x_ref||.  The oracle's own fp32-vs-fp64 spread on these problems is
# ~3e-5 (tests/test_oracle.py), i.e. the tolerance sits just above fp32 round-off.
RTOL = 1e-4


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def run_gpu(p, iters, lm, ep, motion_only=False):
    import droid_backends
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    poses, disps = t(p.poses), t(p.disps)
    dx, dz = droid_backends.ba(poses, disps, t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta),
                               t(p.ii), t(p.jj), p.t0, p.t1, iters, lm, ep, motion_only)
    torch.cuda.synchronize()
    return poses.cpu().numpy(), disps.cpu().numpy(), dx.cpu().numpy(), None if dz is None else dz.cpu().numpy()


def run_oracle(p, iters, lm, ep, motion_only=False):
    from oracle import oracle
    return oracle.ba(p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, p.jj, p.t0, p.t1,
                     iters, lm, ep, motion_only=motion_only)


def check_chained(p, iters, lm, ep, motion_only=False, rtol=RTOL):
    """per-iteration parity: every iteration starts from the ORACLE's state, so round-off of earlier
    iterations (which validity flips at the MIN_DEPTH plane can amplify) does not enter the comparison."""
    import copy
    q = copy.copy(p)
    floor = None
    for _ in range(iters):
        _, _, n0 = check(q, 1, lm, ep, motion_only, rtol, floor)
        floor = floor or 1e-2 * n0
        ref = run_oracle(q, 1, lm, ep, motion_only)
        q = copy.copy(q)
        q.poses, q.disps = ref["poses"], ref["disps"]


def check(p, iters, lm, ep, motion_only=False, rtol=RTOL, floor=None):
    """GPU vs oracle(fp32).  Tolerance: the north_star's 1e-4 normwise, widened only where the reference's own
    fp32 arithmetic is less reproducible than that: tol = max(rtol, 2 * ||oracle_f32 - oracle_f64|| / ||oracle_f64||).
    A result within rtol of the fp64 oracle (the same algorithm without rounding noise) also passes: the kernel's partial
    sums meet in fp64, so it can sit closer to that than the fp32 oracle does, and the distance between two fp32-level
    results is then mostly the oracle's own error.
    `floor` (absolute) keeps the relative error meaningful once dx has converged to round-off level."""
    from oracle import oracle
    poses, disps, dx, dz = run_gpu(p, iters, lm, ep, motion_only)
    ref = run_oracle(p, iters, lm, ep, motion_only)
    r64 = oracle.ba(p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, p.jj, p.t0, p.t1,
                    iters, lm, ep, motion_only=motion_only, dtype=np.float64)
    fl = floor or 0.0

    def relf(a, b):
        a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
        return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), fl, 1e-30))

    tol_x = max(rtol, 2 * relf(ref["dx"], r64["dx"]))
    assert relf(dx, ref["dx"]) <= tol_x or relf(dx, r64["dx"]) <= rtol, ("dx", relf(dx, ref["dx"]), tol_x, relf(dx, r64["dx"]))
    move = np.abs(ref["poses"] - p.poses).max()
    assert np.abs(poses - ref["poses"]).max() <= tol_x * max(move, fl) + 1e-6
    if not motion_only:
        tol_z = max(rtol, 2 * relf(ref["dz"], r64["dz"]))
        assert relf(dz, ref["dz"]) <= tol_z or relf(dz, r64["dz"]) <= rtol, ("dz", relf(dz, ref["dz"]), tol_z, relf(dz, r64["dz"]))
        assert np.abs(disps - ref["disps"]).max() <= tol_z * np.abs(ref["disps"] - p.disps).max() + 1e-6
    else:
        np.testing.assert_array_equal(disps, p.disps)
    return poses, disps, float(np.linalg.norm(ref["dx"]))


def test_ba_c1():
    from droid_b200 import synth
    check_chained(synth.config_c1(), 3, 1e-4, 0.1)
    check(synth.config_c1(), 2, 1e-4, 0.1)          # the in-call iteration loop (itrs=2, factor_graph.py:240)


def test_ba_c1_stereo_sensor_invalid():
    """stereo ii==jj edges, sensor depth prior, points behind the MIN_DEPTH plane"""
    from droid_b200 import synth
    ii, jj = synth.neighborhood_edges(0, 8, 3)
    ii = np.concatenate([ii, np.arange(8)]); jj = np.concatenate([jj, np.arange(8)])
    p = synth.make_ba_problem(ii, jj, 8, seed=77, sensor_frac=0.5, invalid_frac=0.05)
    check_chained(p, 2, 1e-4, 0.1)
    check(p, 2, 1e-4, 0.1)


def test_ba_c2_window():
    from droid_b200 import synth
    check_chained(synth.config_c2(), 2, 1e-4, 0.1)
    check(synth.config_c2(), 2, 1e-4, 0.1)


def test_ba_small_maps_ragged():
    """ht*wd not a multiple of the 256-pixel sub-tile, t0 > 1, buffer larger than t1"""
    from droid_b200 import synth
    rng = np.random.default_rng(3)
    ii, jj = synth.backend_edges(12, 70, rng, rad=2)
    p = synth.make_ba_problem(ii, jj, 12, ht=21, wd=29, seed=5, buffer=20, t0=2)
    check_chained(p, 2, 1e-5, 1e-2)


def test_ba_motion_only():
    """trajectory-filler shape: new frames [t0,t1) with 2 edges each to fixed keyframes (trajectory_filler.py:68-72)"""
    from droid_b200 import synth
    n_kf, n_new = 6, 8
    ii = np.repeat(np.arange(n_kf, n_kf + n_new), 2)
    jj = np.stack([np.arange(n_new) % (n_kf - 1), np.arange(n_new) % (n_kf - 1) + 1], 1).reshape(-1)
    p = synth.make_ba_problem(ii, jj, n_kf + n_new, seed=9, t0=n_kf, t1=n_kf + n_new)
    check_chained(p, 3, 1e-4, 0.1, motion_only=True)


def test_ba_degenerate_calls():
    """what a young map gives: ONE optimised pose (2 keyframes, t0 = 1); an EMPTY edge list (depths move by the sensor prior
    only, dx = 0: kx = arange(t0, t1), droid_kernels.cu:1336-1344); zero iterations (nothing changes, outputs are the zeros of
    the reference's torch::zeros, droid_kernels.cu:1350-1355)"""
    import droid_backends
    from droid_b200 import synth
    from oracle import oracle
    p = synth.make_ba_problem(np.array([0, 1]), np.array([1, 0]), 2, seed=3, t0=1)
    check(p, 2, 1e-4, 0.1)
    # empty edge list, sensor prior on every pixel
    q = synth.make_ba_problem(np.array([0, 1]), np.array([1, 0]), 4, seed=4, t0=1, sensor_frac=1.0)
    e = np.zeros(0, np.int64)
    z = np.zeros((0, 2) + q.disps.shape[1:], np.float32)
    eta = np.full((q.t1 - q.t0,) + q.disps.shape[1:], 1e-3, np.float32)
    ref = oracle.ba(q.poses, q.disps, q.intrinsics, q.disps_sens, z, z, eta, e, e, q.t0, q.t1, 1, 1e-4, 0.1)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    poses, disps = t(q.poses), t(q.disps)
    dx, dz = droid_backends.ba(poses, disps, t(q.intrinsics), t(q.disps_sens), t(z), t(z), t(eta), t(e), t(e), q.t0, q.t1, 1, 1e-4, 0.1, False)
    assert float(dx.abs().max()) == 0.0 and np.array_equal(poses.cpu().numpy(), q.poses)
    assert np.abs(ref["dz"]).max() > 0 and rel(dz.cpu().numpy(), ref["dz"]) <= RTOL
    assert rel(disps.cpu().numpy(), ref["disps"]) <= 1e-6
    # zero iterations
    poses, disps = t(p.poses), t(p.disps)
    dx, dz = droid_backends.ba(poses, disps, t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta), t(p.ii), t(p.jj),
                               p.t0, p.t1, 0, 1e-4, 0.1, False)
    assert np.array_equal(poses.cpu().numpy(), p.poses) and np.array_equal(disps.cpu().numpy(), p.disps)
    assert float(dx.abs().max()) == 0.0 and float(dz.abs().max()) == 0.0


def test_ba_cholesky_failure_gives_zero_pose_update():
    from droid_b200 import synth
    import droid_backends
    p = synth.config_c1()
    ref = run_oracle(p, 1, 0.0, -1e9)
    assert ref["status"] == 1
    poses, disps, dx, dz = run_gpu(p, 1, 0.0, -1e9)
    assert np.all(dx == 0.0)
    np.testing.assert_array_equal(poses, p.poses)
    assert rel(dz, ref["dz"]) <= RTOL


def test_ba_bad_index_and_eta_rows_raise():
    from droid_b200 import synth
    import droid_backends
    p = synth.config_c1()
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    jj = p.jj.copy(); jj[0] = 99
    with pytest.raises(RuntimeError):
        droid_backends.ba(t(p.poses), t(p.disps), t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta),
                          t(p.ii), t(jj), p.t0, p.t1, 1, 1e-4, 0.1, False)
    with pytest.raises(RuntimeError):
        droid_backends.ba(t(p.poses), t(p.disps), t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta[:-1]),
                          t(p.ii), t(p.jj), p.t0, p.t1, 1, 1e-4, 0.1, False)
    with pytest.raises(RuntimeError):
        droid_backends.ba(t(p.poses), t(p.disps), t(p.intrinsics), t(p.disps_sens), t(p.target).permute(0, 1, 3, 2),
                          t(p.weight), t(p.eta), t(p.ii), t(p.jj), p.t0, p.t1, 1, 1e-4, 0.1, False)


def test_ba_ate_within_1pct_of_reference():
    """north_star: final ATE within 1% of the reference on the same synthetic trajectory"""
    from droid_b200 import synth
    p = synth.config_c2()
    poses, _, _, _ = run_gpu(p, 6, 1e-4, 0.1)
    ref = run_oracle(p, 6, 1e-4, 0.1)
    n = p.t1
    gt = synth.camera_centres(p.poses_gt[:n])
    a = synth.umeyama_ate(synth.camera_centres(poses[:n]), gt)
    b = synth.umeyama_ate(synth.camera_centres(ref["poses"][:n]), gt)
    a0 = synth.umeyama_ate(synth.camera_centres(p.poses[:n]), gt)
    assert b < 0.5 * a0            # the optimisation actually converges
    assert abs(a - b) <= 0.01 * b


def test_ba_c3_properties_full_size():
    """BASELINE config 3 (512 KF, ~6k edges): too large for the scalar oracle in a test, so check
    size-independent properties: finite update, cost decrease, fixed pose untouched, determinism to fp32 round-off."""
    from droid_b200 import synth
    p = synth.config_c3()
    poses, disps, dx, dz = run_gpu(p, 2, 1e-5, 1e-2)
    assert np.isfinite(dx).all() and np.isfinite(dz).all()
    np.testing.assert_array_equal(poses[:p.t0], p.poses[:p.t0])
    n = p.t1
    gt = synth.camera_centres(p.poses_gt[:n])
    assert synth.umeyama_ate(synth.camera_centres(poses[:n]), gt) < 0.5 * synth.umeyama_ate(synth.camera_centres(p.poses[:n]), gt)
    poses2, disps2, dx2, dz2 = run_gpu(p, 2, 1e-5, 1e-2)
    assert rel(dx2, dx) < 1e-6 and rel(dz2, dz) < 1e-6


@pytest.mark.parametrize("n_kf", [26, 27, 40, 120, 260])
def test_ba_solver_regimes_parity_and_repeatability(n_kf):
    """The solver changes its schedule with the size of the reduced system: up to 9 tiles per side the round-1 ownership
    (no diagonal CTAs), above that diagonal CTAs + staged tensor-core trailing updates; throughput-bound early steps and a
    chain-bound tail.  Mid-size graphs on small images (the oracle finishes in seconds) cover 4, 12 and 25 tiles per side:
    parity with the oracle, and the same update on every repetition.  26 / 27 keyframes straddle the switch between the
    one-CTA shared-memory solver (n = 150 <= 152, its largest case) and the tiled solver (n = 156)."""
    from droid_b200 import synth
    p = synth.config_c3(seed=2000 + n_kf, n_kf=n_kf, n_edges=8 * n_kf, ht=16, wd=24)
    # 16x24 maps: 8x fewer pixels behind every pose than at 48x64 and hundreds of poses — the fp32 oracle's own dx is 1e-3
    # away from the fp64 one at 260 keyframes and the kernel's dz sits at 1.0e-4 +- 5 % depending on the summation
    # grouping.  This test is about the solver's schedules (a wrong tile shows as 1e-3 .. 1e-1); the 1e-4 claim is
    # carried by the full-resolution tests (c1, c2, test_ba_backend_graph_full_resolution_parity).
    check(p, 1, 1e-5, 1e-2, rtol=2e-4)
    first = None
    for _ in range(6):
        _, _, dx, _ = run_gpu(p, 1, 1e-5, 1e-2)
        if first is None:
            first = dx
        else:
            assert rel(dx, first) < 1e-7


@pytest.mark.parametrize("n_kf", [700, 1024])
def test_ba_solver_intermediate_schedules(n_kf):
    """Between c3 (48 tiles per side: 16 diagonal CTAs) and c4 (144 tiles: 8) the number of diagonal CTAs follows 8 G / nt:
    700 keyframes -> 66 tiles, 16 diagonal CTAs; a full 1024-frame buffer -> 96 tiles, 12.  The solve stage alone (no oracle
    run at this size): the GPU's dx against an fp64 solve of the GPU's own reduced system, and the same update twice."""
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from ba_system import check_solve
    from droid_b200 import synth
    p = synth.config_c3(seed=3000 + n_kf, n_kf=n_kf, n_edges=8 * n_kf, ht=16, wd=24)
    e = check_solve(p, 1e-5, 1e-2)
    print(f"{n_kf} keyframes: dx vs fp64 solve of the GPU system {e:.1e}")
    _, _, dx1, _ = run_gpu(p, 1, 1e-5, 1e-2)
    _, _, dx2, _ = run_gpu(p, 1, 1e-5, 1e-2)
    assert np.isfinite(dx1).all() and rel(dx2, dx1) < 1e-7


def test_ba_backend_graph_full_resolution_parity():
    """Backend-like graph at the full 48x64 map size with hundreds of edges per GPU wave — the regime of c3, where the
    accumulate stage uses few splits per frame and its fp32 chains are longest.  Two precision bugs lived here unseen while
    c3 was only checked for properties: the tensor core's truncating accumulator addition (dx 2e-3 off) and 24-chunk fp32
    chains (1.9e-3 off on this graph).  The graph is ill-conditioned enough that the reference's own fp32 arithmetic is 1e-3
    away from exact arithmetic in dx, and ANY fp32 grouping of the sums scatters by that much (accumulate splits 2..24:
    2e-4 .. 1.9e-3), so the claim checked is the same as everywhere: within max(1e-4, 2 x the fp32 oracle's own distance
    from the fp64 oracle) — here also measured against the fp64 oracle directly."""
    from droid_b200 import synth
    from oracle import oracle
    p = synth.config_c3(seed=4242, n_kf=160, n_edges=1600)
    check(p, 1, 1e-5, 1e-2)
    _, _, dx, dz = run_gpu(p, 1, 1e-5, 1e-2)
    a = (p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, p.jj, p.t0, p.t1, 1, 1e-5, 1e-2)
    r32, r64 = oracle.ba(*a), oracle.ba(*a, dtype=np.float64)
    assert rel(dx, r64["dx"]) <= max(RTOL, 2 * rel(r32["dx"], r64["dx"])), (rel(dx, r64["dx"]), rel(r32["dx"], r64["dx"]))
    assert rel(dz, r64["dz"]) <= max(RTOL, 2 * rel(r32["dz"], r64["dz"])), (rel(dz, r64["dz"]), rel(r32["dz"], r64["dz"]))


def test_ba_c3_repeatable():
    """Race detector for the solver's dataflow (static tile ownership, flags, streamed hand-overs, staged operands): the
    same call repeated must give the same update.  The only run-to-run freedom is the order of the fp64 atomics of the
    Schur scatter (1e-13 .. 4e-10 on dx at this size, a handful of discrete values); a tile read before it is final shows
    up as 1e-4 .. 1e-1."""
    from droid_b200 import synth
    p = synth.config_c3()
    first = None
    for _ in range(12):
        _, _, dx, _ = run_gpu(p, 2, 1e-5, 1e-2)
        if first is None:
            first = dx
        else:
            assert rel(dx, first) < 1e-8


@pytest.mark.timeout(600)
def test_ba_c4_multisession_size():
    """BASELINE config 4 (3 fused maps, 1536 keyframes, ~20k edges incl. stereo and cross-map edges) on ONE GPU:
    a 9210 x 9210 reduced system (144 solver tiles), 1536 depth maps.  Size-independent properties only."""
    import time
    from droid_b200 import synth
    p = synth.config_c4()
    assert p.t1 - p.t0 >= 1500 and p.n_edges >= 19000
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    import droid_backends
    args = [t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta), t(p.ii), t(p.jj)]
    poses, disps = t(p.poses), t(p.disps)
    dx, dz = droid_backends.ba(poses, disps, *args, p.t0, p.t1, 2, 1e-5, 1e-2, False)
    torch.cuda.synchronize()
    poses2, disps2 = t(p.poses), t(p.disps)
    t0 = time.perf_counter()
    dx2, dz2 = droid_backends.ba(poses2, disps2, *args, p.t0, p.t1, 2, 1e-5, 1e-2, False)
    torch.cuda.synchronize()
    print(f"c4 ba: {p.n_edges} edges, {p.t1 - p.t0} poses, 2 iterations: {1e3 * (time.perf_counter() - t0):.1f} ms")
    dxn, dzn = dx.cpu().numpy(), dz.cpu().numpy()
    assert np.isfinite(dxn).all() and np.isfinite(dzn).all()
    assert rel(dx2.cpu().numpy(), dxn) < 1e-6
    np.testing.assert_array_equal(poses.cpu().numpy()[:p.t0], p.poses[:p.t0])
    n = p.t1
    gt = synth.camera_centres(p.poses_gt[:n])
    before = synth.umeyama_ate(synth.camera_centres(p.poses[:n]), gt)
    after = synth.umeyama_ate(synth.camera_centres(poses.cpu().numpy()[:n]), gt)
    assert after < 0.7 * before                 # two Gauss-Newton steps: measured 0.0175 -> 0.0094


@pytest.mark.parametrize("beta", [0.3, 0.25])
def test_frame_distance_bit_exact(beta):
    import droid_backends
    from oracle import oracle
    from droid_b200 import synth
    p = synth.config_c2()
    n = p.t1
    ii, jj = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    ii, jj = ii.reshape(-1).astype(np.int64), jj.reshape(-1).astype(np.int64)
    disps = p.disps.copy()
    disps[3] = 1e3     # frame 3's points land behind / within MIN_DEPTH of most other cameras -> 1000.0 branch
    ref = oracle.frame_distance(p.poses, disps, p.intrinsics, ii, jj, beta)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    out = droid_backends.frame_distance(t(p.poses), t(disps), t(p.intrinsics), t(ii), t(jj), beta).cpu().numpy()
    assert (ref == 1000.0).any() and (ref != 1000.0).any()
    np.testing.assert_array_equal(out.view(np.uint32), ref.view(np.uint32))
    # the consumer's argsort (factor_graph.py:354) is therefore identical
    np.testing.assert_array_equal(np.argsort(out, kind="stable"), np.argsort(ref, kind="stable"))


@pytest.mark.parametrize("ht,wd", [(48, 64), (40, 64), (24, 32)])
def test_frame_distance_batched_kernel_bit_exact(ht, wd):
    """Enough pairs (>= 4736) to take the batched kernel (8 pairs per CTA, pair-independent work hoisted): still every
    bit of the oracle, for the 48x64 and the fork's 40x64 map sizes, ragged pair count, pairs that change their source
    frame inside a batch, an invalid index and the 1000.0 branch."""
    import droid_backends
    from oracle import oracle
    from droid_b200 import synth
    n = 72
    ii0, jj0 = synth.neighborhood_edges(0, n, 2)
    p = synth.make_ba_problem(ii0, jj0, n, ht=ht, wd=wd, seed=31)
    ii, jj = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    ii, jj = ii.reshape(-1).astype(np.int64), jj.reshape(-1).astype(np.int64)
    rng = np.random.default_rng(2)
    perm = rng.permutation(len(ii))[:1003]                       # + a ragged tail of pairs in random order
    ii, jj = np.concatenate([ii, ii[perm]]), np.concatenate([jj, jj[perm]])
    assert len(ii) >= 4736 and len(ii) % 8 != 0
    disps = p.disps.copy(); disps[5] = 1e3
    ref = oracle.frame_distance(p.poses, disps, p.intrinsics, ii, jj, 0.25)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    out = droid_backends.frame_distance(t(p.poses), t(disps), t(p.intrinsics), t(ii), t(jj), 0.25).cpu().numpy()
    assert (ref == 1000.0).any() and (ref != 1000.0).any()
    np.testing.assert_array_equal(out.view(np.uint32), ref.view(np.uint32))
    ii[17] = n + 3                                               # out of range: NaN for that pair only
    out2 = droid_backends.frame_distance(t(p.poses), t(disps), t(p.intrinsics), t(ii), t(jj), 0.25).cpu().numpy()
    assert np.isnan(out2[17]) and np.array_equal(np.delete(out2, 17).view(np.uint32), np.delete(ref, 17).view(np.uint32))


def test_iproj_depth_filter():
    import droid_backends
    from oracle import oracle
    from droid_b200 import synth
    p = synth.config_c1()
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    pts = droid_backends.iproj(t(p.poses), t(p.disps), t(p.intrinsics)).cpu().numpy()
    np.testing.assert_allclose(pts, oracle.iproj(p.poses, p.disps, p.intrinsics), rtol=1e-4, atol=1e-5)
    inds = np.arange(8, dtype=np.int64); th = np.full(8, 0.05, np.float32)
    cnt = droid_backends.depth_filter(t(p.poses_gt.astype(np.float32)), t(p.disps_gt.astype(np.float32)), t(p.intrinsics),
                                      t(inds), t(th)).cpu().numpy()
    ref = oracle.depth_filter(p.poses_gt.astype(np.float32), p.disps_gt.astype(np.float32), p.intrinsics, inds, th)
    assert (cnt != ref).mean() < 1e-3 and ref.max() >= 1
