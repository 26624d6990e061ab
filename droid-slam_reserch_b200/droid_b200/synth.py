"""Synthetic EuRoC-shaped inputs for the BA / correlation hot path.

Everything is generated with numpy from a fixed seed so that tests, golden-vector scripts, the
oracle and bench.py all see bit-identical inputs.  Shapes and conventions follow the reference:
poses are world->camera `(tx,ty,tz,qx,qy,qz,qw)` (droid_slam/depth_video.py:33), disparities are
inverse depths at 1/8 resolution, intrinsics are EuRoC's rectified pinhole scaled to the working
image and divided by 8 (evaluation_scripts/test_euroc.py:51,70-74, droid_slam/motion_filter.py:78),
`target`/`weight` are `[N,2,ht,wd]` (droid_slam/factor_graph.py:237-238) and `eta` is
`.2*damping[unique(ii)] + EP` (factor_graph.py:234).
"""
from __future__ import annotations

import dataclasses
import numpy as np

EUROC_FX, EUROC_FY, EUROC_CX, EUROC_CY = 435.2, 435.2, 367.45, 252.2  # on 752x480 rectified images


def intrinsics_eighth(image_hw=(384, 512)) -> np.ndarray:
    h, w = image_hw
    sx, sy = w / 752.0, h / 480.0
    return (np.array([EUROC_FX * sx, EUROC_FY * sy, EUROC_CX * sx, EUROC_CY * sy]) / 8.0).astype(np.float32)


# ----------------------------------------------------------------------------- SE3 (float64, generation only)
def quat_mul(a, b):
    ax, ay, az, aw = a[..., 0], a[..., 1], a[..., 2], a[..., 3]
    bx, by, bz, bw = b[..., 0], b[..., 1], b[..., 2], b[..., 3]
    return np.stack([aw * bx + ax * bw + ay * bz - az * by,
                     aw * by - ax * bz + ay * bw + az * bx,
                     aw * bz + ax * by - ay * bx + az * bw,
                     aw * bw - ax * bx - ay * by - az * bz], -1)


def quat_rot(q, v):
    qv = q[..., :3]
    uv = 2.0 * np.cross(qv, v)
    return v + q[..., 3:4] * uv + np.cross(qv, uv)


def se3_exp(xi):
    xi = np.asarray(xi, np.float64)
    tau, phi = xi[..., :3], xi[..., 3:]
    th = np.linalg.norm(phi, axis=-1, keepdims=True)
    small = th < 1e-6
    ths = np.where(small, 1.0, th)
    imag = np.where(small, 0.5 - th ** 2 / 48.0, np.sin(0.5 * ths) / ths)
    q = np.concatenate([imag * phi, np.cos(0.5 * th)], -1)
    a = np.where(small, 0.5, (1 - np.cos(ths)) / ths ** 2)
    b = np.where(small, 1.0 / 6.0, (ths - np.sin(ths)) / ths ** 3)
    c1 = np.cross(phi, tau)
    c2 = np.cross(phi, c1)
    return np.concatenate([tau + a * c1 + b * c2, q], -1)


def se3_mul(a, b):
    return np.concatenate([a[..., :3] + quat_rot(a[..., 3:], b[..., :3]), quat_mul(a[..., 3:], b[..., 3:])], -1)


def se3_inv(a):
    qi = a[..., 3:] * np.array([-1.0, -1.0, -1.0, 1.0])
    return np.concatenate([-quat_rot(qi, a[..., :3]), qi], -1)


def helix_trajectory(n: int, rng: np.random.Generator) -> np.ndarray:
    """Smooth helix; 0.05-0.15 m and 1-3 deg between keyframes.  Returns world->camera poses [n,7] f64."""
    step = rng.uniform(0.05, 0.15, n)
    dang = np.deg2rad(rng.uniform(1.0, 3.0, n))
    s = np.cumsum(step)
    ang = np.cumsum(dang)
    radius = 2.0
    # camera centres on a helix, looking roughly along the tangent with a slow roll
    c = np.stack([radius * np.cos(s / radius), radius * np.sin(s / radius), 0.05 * s], -1)
    poses = np.zeros((n, 7))
    for k in range(n):
        yaw = se3_exp(np.array([0, 0, 0, 0, ang[k], 0.0]))
        roll = se3_exp(np.array([0, 0, 0, 0, 0, 0.2 * np.sin(0.1 * k)]))
        q = quat_mul(yaw[3:], roll[3:])  # camera->world rotation
        Twc = np.concatenate([c[k], q])
        poses[k] = se3_inv(Twc)
    return poses


def smooth_disps(n: int, ht: int, wd: int, rng: np.random.Generator, lo=0.2, hi=2.0) -> np.ndarray:
    """Piecewise-planar inverse-depth maps in [lo,hi]."""
    yy, xx = np.meshgrid(np.linspace(-1, 1, ht), np.linspace(-1, 1, wd), indexing="ij")
    out = np.empty((n, ht, wd))
    for k in range(n):
        base = rng.uniform(lo + 0.3, hi - 0.3)
        gx, gy = rng.uniform(-0.25, 0.25, 2)
        d = base + gx * xx + gy * yy
        # a few rectangular patches at other depths
        for _ in range(3):
            y0, x0 = rng.integers(0, ht - 8), rng.integers(0, wd - 8)
            y1, x1 = y0 + rng.integers(6, ht // 2), x0 + rng.integers(6, wd // 2)
            d[y0:y1, x0:x1] += rng.uniform(-0.3, 0.3)
        out[k] = np.clip(d, lo, hi)
    return out


def reproject(poses, disps, intr, ii, jj, stereo_baseline=(-0.1, 0.0, 0.0)):
    """coords[N,ht,wd,2] and depth z of points of frame ii seen from jj (f64).  ii==jj -> fixed stereo
    baseline (droid_kernels.cu:219-229)."""
    fx, fy, cx, cy = [float(v) for v in intr]
    n, ht, wd = disps.shape
    y, x = np.meshgrid(np.arange(ht, dtype=np.float64), np.arange(wd, dtype=np.float64), indexing="ij")
    X = np.stack([(x - cx) / fx, (y - cy) / fy, np.ones_like(x)], -1)
    out = np.empty((len(ii), ht, wd, 2))
    zs = np.empty((len(ii), ht, wd))
    for e, (i, j) in enumerate(zip(ii, jj)):
        if i == j:
            G = np.array([*stereo_baseline, 0, 0, 0, 1.0])
        else:
            G = se3_mul(poses[j], se3_inv(poses[i]))
        Y = quat_rot(G[3:], X) + disps[i][..., None] * G[:3]
        z = Y[..., 2]
        out[e, ..., 0] = fx * Y[..., 0] / z + cx
        out[e, ..., 1] = fy * Y[..., 1] / z + cy
        zs[e] = z
    return out, zs


# ----------------------------------------------------------------------------- graphs
def neighborhood_edges(t0: int, t1: int, r: int = 3, stereo: bool = False):
    """FactorGraph.add_neighborhood_factors pattern (factor_graph.py:302-312)."""
    c = 1 if stereo else 0
    ii, jj = np.meshgrid(np.arange(t0, t1), np.arange(t0, t1), indexing="ij")
    ii, jj = ii.reshape(-1), jj.reshape(-1)
    keep = (np.abs(ii - jj) > c) & (np.abs(ii - jj) <= r)
    return ii[keep].astype(np.int64), jj[keep].astype(np.int64)


def backend_edges(n_kf: int, n_edges: int, rng: np.random.Generator, rad: int = 2, stereo_frac: float = 0.0,
                  loop_span: int = 0):
    """Global-backend style graph (factor_graph.py:315-379): bidirectional temporal edges i<->j for
    i-rad-1 <= j < i, optional stereo self edges, then random bidirectional "proximity" edges until
    about `n_edges` edges exist.  Proximity partners are drawn near revisited places: frame i is paired
    with i - k*period +- jitter so that the reduced system has realistic loop-closure structure."""
    es = []
    seen = set()

    def add(i, j):
        if (i, j) not in seen:
            seen.add((i, j)); es.append((i, j))

    n_st = int(round(stereo_frac * n_kf))
    st_frames = set(rng.choice(n_kf, n_st, replace=False).tolist()) if n_st else set()
    for i in range(n_kf):
        if i in st_frames:
            add(i, i)
        for j in range(max(i - rad - 1, 0), i):
            add(i, j); add(j, i)
    period = loop_span if loop_span > 0 else max(n_kf // 4, 8)
    guard = 0
    while len(es) < n_edges and guard < 50 * n_edges:
        guard += 1
        i = int(rng.integers(rad + 2, n_kf))
        if rng.random() < 0.6:
            j = i - int(rng.integers(rad + 2, min(i, 24) + 1)) if i > rad + 2 else -1
        else:
            k = int(rng.integers(1, max(i // period, 1) + 1))
            j = i - k * period + int(rng.integers(-6, 7))
        if j < 0 or j >= n_kf or abs(i - j) <= rad + 1:
            continue
        add(i, j); add(j, i)
    ii = np.array([e[0] for e in es], np.int64)
    jj = np.array([e[1] for e in es], np.int64)
    return ii, jj


def hub_edges(n_kf: int, hubs, rng: np.random.Generator, rad: int = 2):
    """Temporal graph (i <-> j for i-rad-1 <= j < i) + hub frames: `hubs` = {frame: out-degree}; hub h gets edges h -> j
    and j -> h to `out-degree` distinct other frames in total (its temporal neighbours count) — the shape of a place that
    is revisited many times in a global / multisession backend graph (droid_backend.py:32-39 asks for 16 t factors)."""
    es, seen = [], set()

    def add(i, j):
        if i != j and (i, j) not in seen:
            seen.add((i, j)); es.append((i, j))

    for i in range(n_kf):
        for j in range(max(i - rad - 1, 0), i):
            add(i, j); add(j, i)
    for h, deg in hubs.items():
        have = {j for (i, j) in seen if i == h}
        others = [j for j in rng.permutation(n_kf).tolist() if j not in hubs and j not in have]   # hubs keep their exact degree
        for j in others[:max(deg - len(have), 0)]:
            add(h, j); add(j, h)
    ii = np.array([e[0] for e in es], np.int64)
    jj = np.array([e[1] for e in es], np.int64)
    return ii, jj


def hot_place_trajectory(visits: int = 12, arc: int = 8, away: int = 8, radius: float = 1.5):
    """A camera that walks the same `arc` frames of a circle (looking outward) on every one of `visits` visits and spends
    `away` frames on an excursion of its own in between: the shape of a multisession / long-term map with one place seen
    again and again.  `add_proximity_factors` with the backend's budget (max_factors = 16 t, droid_backend.py:32-39) then
    concentrates its loop-closure edges on the hot frames: 12 visits -> out-degrees up to 36, 18 frames above 28.
    Returns world->camera poses [visits * (arc + away), 7] f64."""
    poses = []
    for v in range(visits):
        for k in range(arc + away):
            if k < arc:
                a = 0.10 * k + 0.004 * v
                c = np.array([radius * np.cos(a), 0.01 * v, radius * np.sin(a)])
                yaw = np.pi / 2 - a
            else:
                u = k - arc + 1
                a = 0.10 * (arc - 1)
                base = np.array([radius * np.cos(a), 0.0, radius * np.sin(a)])
                ang = 2 * np.pi * v / visits
                c = base + 0.5 * u * np.array([np.cos(ang), 0.3, np.sin(ang)]) + np.array([0, 2.0 + 0.8 * v, 0]) * min(u, 2) / 2
                yaw = np.pi / 2 - a + 0.5 * u * (1 if v % 2 else -1)
            q = se3_exp(np.array([0, 0, 0, 0, yaw, 0.0]))[3:]
            poses.append(se3_inv(np.concatenate([c, q])))
    return np.array(poses)


# ----------------------------------------------------------------------------- BA problem
@dataclasses.dataclass
class BAProblem:
    poses: np.ndarray        # [B,7] f32 initial (perturbed) world->camera
    disps: np.ndarray        # [B,ht,wd] f32 initial
    intrinsics: np.ndarray   # [4] f32
    disps_sens: np.ndarray   # [B,ht,wd] f32 (0 = no sensor depth)
    target: np.ndarray       # [N,2,ht,wd] f32
    weight: np.ndarray       # [N,2,ht,wd] f32
    eta: np.ndarray          # [M,ht,wd] f32
    ii: np.ndarray           # [N] i64
    jj: np.ndarray           # [N] i64
    t0: int
    t1: int
    poses_gt: np.ndarray     # [B,7] f64
    disps_gt: np.ndarray     # [B,ht,wd] f64

    @property
    def n_edges(self):
        return int(self.ii.shape[0])


def make_ba_problem(ii, jj, n_kf, ht=48, wd=64, seed=0, buffer=None, t0=1, t1=None,
                    pose_sigma=0.01, disp_sigma=0.05, pix_sigma=0.5, invalid_frac=0.0,
                    sensor_frac=0.0, intr=None, poses_gt=None, disps_gt=None) -> BAProblem:
    rng = np.random.default_rng(seed)
    B = buffer or n_kf
    t1 = n_kf if t1 is None else t1
    intr = intrinsics_eighth((ht * 8, wd * 8)) if intr is None else np.asarray(intr, np.float32)
    given_p, given_d = poses_gt, disps_gt
    poses_gt = np.tile(np.array([0, 0, 0, 0, 0, 0, 1.0]), (B, 1))
    traj = helix_trajectory(n_kf, rng)                      # always drawn: the random stream does not depend on the options
    poses_gt[:n_kf] = traj if given_p is None else np.asarray(given_p, np.float64)[:n_kf]
    disps_gt = np.ones((B, ht, wd))
    sm_d = smooth_disps(n_kf, ht, wd, rng)
    disps_gt[:n_kf] = sm_d if given_d is None else np.asarray(given_d, np.float64)[:n_kf]
    if invalid_frac > 0:  # make some points land behind / too close to the target camera
        mask = rng.random((n_kf, ht, wd)) < invalid_frac
        disps_gt[:n_kf][mask] = 40.0
    coords, _ = reproject(poses_gt, disps_gt, intr, ii, jj)
    N = len(ii)
    target = coords.transpose(0, 3, 1, 2) + rng.normal(0, pix_sigma, (N, 2, ht, wd))
    weight = rng.uniform(0, 1, (N, 2, ht, wd))
    weight[rng.random((N, 2, ht, wd)) < 0.1] = 0.0
    # initial estimate: GT perturbed on the left (same side as the retraction, droid_kernels.cu:877-895)
    poses = poses_gt.copy()
    noise = rng.normal(0, pose_sigma, (n_kf, 6))
    noise[:t0] = 0.0
    poses[:n_kf] = se3_mul(se3_exp(noise), poses_gt[:n_kf])
    disps = disps_gt.copy()
    disps[:n_kf] *= np.exp(rng.normal(0, disp_sigma, (n_kf, 1, 1))) * (1 + rng.normal(0, 0.02, (n_kf, ht, wd)))
    disps_sens = np.zeros((B, ht, wd))
    if sensor_frac > 0:
        sel = rng.random((n_kf, ht, wd)) < sensor_frac
        disps_sens[:n_kf][sel] = (disps_gt[:n_kf] * (1 + rng.normal(0, 0.01, (n_kf, ht, wd))))[sel]
    kx = np.unique(np.concatenate([np.arange(t0, t1), ii]))
    eta = 0.2 * rng.uniform(1e-3, 1e-1, (len(kx), ht, wd)) + 1e-7
    f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    return BAProblem(f32(poses), f32(disps), intr, f32(disps_sens), f32(target), f32(weight), f32(eta),
                     np.ascontiguousarray(ii, np.int64), np.ascontiguousarray(jj, np.int64), int(t0), int(t1),
                     poses_gt, disps_gt)


def config_c1(seed=1235, stereo=False):
    """8 keyframes, 48x64, radius-3 neighbourhood graph, t0=1 (BASELINE.json configs[0])."""
    ii, jj = neighborhood_edges(0, 8, 3, stereo)
    return make_ba_problem(ii, jj, 8, seed=seed)


def config_c2(seed=1236, n_kf=25, n_edges=100):
    """mono frontend window: ~100 edges over a 25-keyframe window, t0 = window start."""
    rng = np.random.default_rng(seed)
    ii, jj = backend_edges(n_kf, n_edges, rng, rad=2)
    return make_ba_problem(ii, jj, n_kf, seed=seed, buffer=n_kf + 7, t0=3)


def config_c3(seed=1237, n_kf=512, n_edges=6000, stereo_frac=0.0, ht=48, wd=64, sensor_frac=0.0):
    """global backend: 512 keyframes, ~6k edges, t0=1, t1=n_kf (droid_backend.py / factor_graph.py:297)."""
    rng = np.random.default_rng(seed)
    ii, jj = backend_edges(n_kf, n_edges, rng, rad=2, stereo_frac=stereo_frac)
    return make_ba_problem(ii, jj, n_kf, ht=ht, wd=wd, seed=seed, sensor_frac=sensor_frac)


def config_c4(seed=1238, maps=3, kf_per_map=512, edges_per_map=6400, cross_edges=800):
    """multisession: `maps` trajectories concatenated + cross-map proximity edges (BackendAllMaps.py)."""
    rng = np.random.default_rng(seed)
    iis, jjs = [], []
    for m in range(maps):
        a, b = backend_edges(kf_per_map, edges_per_map, rng, rad=2, stereo_frac=1.0)
        iis.append(a + m * kf_per_map); jjs.append(b + m * kf_per_map)
    n_kf = maps * kf_per_map
    seen = set(zip(np.concatenate(iis).tolist(), np.concatenate(jjs).tolist()))
    ci, cj = [], []
    while len(ci) < cross_edges:
        i, j = int(rng.integers(0, n_kf)), int(rng.integers(0, n_kf))
        if i // kf_per_map == j // kf_per_map or (i, j) in seen:
            continue
        seen.add((i, j)); seen.add((j, i))
        ci += [i, j]; cj += [j, i]
    ii = np.concatenate(iis + [np.array(ci, np.int64)])
    jj = np.concatenate(jjs + [np.array(cj, np.int64)])
    return make_ba_problem(ii, jj, n_kf, seed=seed)


def make_fmaps(n, ht=48, wd=64, ch=128, seed=0, dtype=np.float16):
    rng = np.random.default_rng(seed)
    return rng.normal(0, 1, (n, ch, ht, wd)).astype(dtype)


def lookup_coords(prob: BAProblem, jitter=0.25, seed=0):
    """coords[N,ht,wd,2] (x,y): reprojection of ii into jj from the current estimate, so that lookup
    windows cross the image border the way they do in FactorGraph.update (factor_graph.py:203-208)."""
    rng = np.random.default_rng(seed)
    c, _ = reproject(prob.poses.astype(np.float64), prob.disps.astype(np.float64), prob.intrinsics, prob.ii, prob.jj)
    c = np.nan_to_num(c, nan=0.0, posinf=1e4, neginf=-1e4)
    return (c + rng.normal(0, jitter, c.shape)).astype(np.float32)


def umeyama_ate(est_centres: np.ndarray, gt_centres: np.ndarray, with_scale=True) -> float:
    """Translation RMSE after Sim(3)/SE(3) alignment — the evo APE used by the reference's eval
    scripts (evaluation_scripts/test_euroc.py:142-145) restated without evo."""
    x, y = np.asarray(est_centres, np.float64), np.asarray(gt_centres, np.float64)
    mx, my = x.mean(0), y.mean(0)
    xc, yc = x - mx, y - my
    S = yc.T @ xc / len(x)
    U, D, Vt = np.linalg.svd(S)
    W = np.eye(3)
    if np.linalg.det(U) * np.linalg.det(Vt) < 0:
        W[2, 2] = -1
    R = U @ W @ Vt
    s = (D * np.diag(W)).sum() / (xc ** 2).sum() * len(x) if with_scale else 1.0
    t = my - s * R @ mx
    err = (s * (R @ x.T).T + t) - y
    return float(np.sqrt((err ** 2).sum(1).mean()))


def camera_centres(poses_w2c: np.ndarray) -> np.ndarray:
    p = np.asarray(poses_w2c, np.float64)
    return se3_inv(p)[..., :3]
