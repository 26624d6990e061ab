"""ctypes front-end of the CPU oracle (oracle/oracle.cpp).  TEST INFRASTRUCTURE ONLY.

May be imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference`
legs — never by the product package.  All functions take and return numpy arrays with the layouts of
the reference ops (src/droid.cpp:237-250).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_REF = None

_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
_u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")


def build(force: bool = False) -> str:
    """Compile oracle.cpp (and oracle/_ref when /root/reference exists).  Returns the .so path."""
    so = os.path.join(_HERE, "_build", "liboracle.so")
    src = os.path.join(_HERE, "oracle.cpp")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "all"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.oracle_ba_f32.restype = C.c_int
        L.oracle_ba_f64.restype = C.c_int
        L.oracle_num_threads.restype = C.c_int
        _LIB = L
    return _LIB


def num_threads() -> int:
    return int(lib().oracle_num_threads())


def _p(a, t=C.c_void_p):
    return None if a is None else a.ctypes.data_as(t)


def ba(poses, disps, intrinsics, disps_sens, target, weight, eta, ii, jj, t0, t1, iterations, lm, ep,
       motion_only=False, skip_pose0_backsub=True, dtype=np.float32, dump_system=False,
       damp_before_schur=False, moba_block_damping=False):
    """droid_backends.ba restated (src/droid_kernels.cu:1314-1434).  Returns a dict with the updated
    `poses`, `disps` (copies — inputs are not modified), `dx`, `dz`, `status` (1 = Cholesky failed →
    zero pose update, like droid_kernels.cu:1207-1210) and optionally the undamped reduced system."""
    f = np.dtype(dtype)
    cv = lambda a: np.ascontiguousarray(a, dtype=f).copy()
    poses, disps = cv(poses), cv(disps)
    intr, sens, target, weight = cv(intrinsics), cv(disps_sens), cv(target), cv(weight)
    ii = np.ascontiguousarray(ii, np.int64); jj = np.ascontiguousarray(jj, np.int64)
    N = int(ii.shape[0]); B, ht, wd = disps.shape
    P = t1 - t0
    kx = np.unique(np.concatenate([np.arange(t0, t1), ii]))
    M = len(kx)
    eta_a = cv(eta) if eta is not None else np.zeros((M, ht, wd), f)
    dx = np.zeros((P, 6), f); dz = np.zeros((M, ht * wd), f)
    args = [_p(poses), _p(disps), _p(intr), _p(sens), _p(target), _p(weight), _p(eta_a), _p(ii), _p(jj),
            C.c_int(N), C.c_int(B), C.c_int(ht), C.c_int(wd), C.c_int(t0), C.c_int(t1), C.c_int(iterations),
            C.c_float(lm), C.c_float(ep), C.c_int(int(motion_only)), C.c_int(int(eta_a.shape[0])), _p(dx), _p(dz),
            C.c_int(int(skip_pose0_backsub) | (int(damp_before_schur) << 1) | (int(moba_block_damping) << 2))]
    out = {}
    if f == np.float32:
        H = np.zeros((6 * P, 6 * P), np.float64) if dump_system else None
        b = np.zeros(6 * P, np.float64) if dump_system else None
        st = lib().oracle_ba_f32(*args, _p(H), _p(b))
        if dump_system:
            out["H"], out["b"] = H, b
    else:
        st = lib().oracle_ba_f64(*args)
    if st < 0:
        raise RuntimeError(f"oracle ba: invalid arguments (status {st})")
    out.update(poses=poses, disps=disps, dx=dx, dz=None if motion_only else dz, status=st, kx=kx)
    return out


def frame_distance(poses, disps, intrinsics, ii, jj, beta):
    poses = np.ascontiguousarray(poses, np.float32); disps = np.ascontiguousarray(disps, np.float32)
    intr = np.ascontiguousarray(intrinsics, np.float32)
    ii = np.ascontiguousarray(ii, np.int64); jj = np.ascontiguousarray(jj, np.int64)
    K = int(ii.shape[0]); _, ht, wd = disps.shape
    out = np.zeros(K, np.float32)
    lib().oracle_frame_distance(_p(poses), _p(disps), _p(intr), _p(ii), _p(jj), C.c_int(K), C.c_int(ht), C.c_int(wd),
                                C.c_float(beta), _p(out))
    return out


def corr_index_forward(volume, coords, radius):
    """volume[N,h1,w1,h2,w2] (f32 or f16), coords[N,2,h1,w1] f32 -> corr[N,2r+1,2r+1,h1,w1]."""
    N, h1, w1, h2, w2 = volume.shape
    rd = 2 * radius + 1
    coords = np.ascontiguousarray(coords, np.float32)
    dims = (C.c_int(N), C.c_int(h1), C.c_int(w1), C.c_int(h2), C.c_int(w2), C.c_int(radius))
    if volume.dtype == np.float16:
        v = np.ascontiguousarray(volume).view(np.uint16)
        out = np.zeros((N, rd, rd, h1, w1), np.uint16)
        lib().oracle_corr_index_forward_f16(_p(v), _p(coords), _p(out), *dims)
        return out.view(np.float16)
    v = np.ascontiguousarray(volume, np.float32)
    out = np.zeros((N, rd, rd, h1, w1), np.float32)
    lib().oracle_corr_index_forward_f32(_p(v), _p(coords), _p(out), *dims)
    return out


def corr_index_backward(volume_shape, coords, corr_grad, radius):
    N, h1, w1, h2, w2 = volume_shape
    coords = np.ascontiguousarray(coords, np.float32); g = np.ascontiguousarray(corr_grad, np.float32)
    out = np.zeros(volume_shape, np.float32)
    lib().oracle_corr_index_backward_f32(_p(coords), _p(g), _p(out), C.c_int(N), C.c_int(h1), C.c_int(w1), C.c_int(h2),
                                         C.c_int(w2), C.c_int(radius))
    return out


def altcorr_forward(fmap1, fmap2, coords, radius):
    """fmap1[B,H1,W1,C], fmap2[B,H2,W2,C], coords[B,S,H1,W1,2] -> corr[B,S,(2r+1)^2,H1,W1] (f32)."""
    f1 = np.ascontiguousarray(fmap1, np.float32); f2 = np.ascontiguousarray(fmap2, np.float32)
    co = np.ascontiguousarray(coords, np.float32)
    B, H1, W1, Cc = f1.shape; _, H2, W2, _ = f2.shape; S = co.shape[1]
    rd = 2 * radius + 1
    out = np.zeros((B, S, rd * rd, H1, W1), np.float32)
    lib().oracle_altcorr_forward(_p(f1), _p(f2), _p(co), _p(out), C.c_int(B), C.c_int(S), C.c_int(H1), C.c_int(W1),
                                 C.c_int(H2), C.c_int(W2), C.c_int(Cc), C.c_int(radius))
    return out


def altcorr_backward(fmap1, fmap2, coords, corr_grad, radius):
    f1 = np.ascontiguousarray(fmap1, np.float32); f2 = np.ascontiguousarray(fmap2, np.float32)
    co = np.ascontiguousarray(coords, np.float32); g = np.ascontiguousarray(corr_grad, np.float32)
    B, H1, W1, Cc = f1.shape; _, H2, W2, _ = f2.shape; S = co.shape[1]
    g1 = np.zeros_like(f1); g2 = np.zeros_like(f2)
    lib().oracle_altcorr_backward(_p(f1), _p(f2), _p(co), _p(g), _p(g1), _p(g2), C.c_int(B), C.c_int(S), C.c_int(H1),
                                  C.c_int(W1), C.c_int(H2), C.c_int(W2), C.c_int(Cc), C.c_int(radius))
    return g1, g2, np.zeros(co.shape, np.float32)


def corr_volume(fmap1, fmap2, num_levels=4):
    """fmap1,fmap2[N,C,ht,wd] f32 -> list of `num_levels` volumes [N,ht,wd,ht>>l,wd>>l] (corr.py:30-38)."""
    f1 = np.ascontiguousarray(fmap1, np.float32); f2 = np.ascontiguousarray(fmap2, np.float32)
    N, Cc, ht, wd = f1.shape
    lv = [np.zeros((N, ht, wd, ht >> l, wd >> l), np.float32) for l in range(num_levels)]
    ptrs = [_p(lv[l]) if l < num_levels else None for l in range(4)]
    lib().oracle_corr_volume_f32(_p(f1), _p(f2), C.c_int(N), C.c_int(Cc), C.c_int(ht), C.c_int(wd), *ptrs)
    return lv


def iproj(poses, disps, intrinsics):
    poses = np.ascontiguousarray(poses, np.float32); disps = np.ascontiguousarray(disps, np.float32)
    num, ht, wd = disps.shape
    out = np.zeros((num, ht, wd, 3), np.float32)
    lib().oracle_iproj(_p(poses), _p(disps), _p(np.ascontiguousarray(intrinsics, np.float32)), C.c_int(num), C.c_int(ht),
                       C.c_int(wd), _p(out))
    return out


def depth_filter(poses, disps, intrinsics, inds, thresh):
    poses = np.ascontiguousarray(poses, np.float32); disps = np.ascontiguousarray(disps, np.float32)
    inds = np.ascontiguousarray(inds, np.int64); thresh = np.ascontiguousarray(thresh, np.float32)
    num, ht, wd = disps.shape
    out = np.zeros((len(inds), ht, wd), np.float32)
    lib().oracle_depth_filter(_p(poses), _p(disps), _p(np.ascontiguousarray(intrinsics, np.float32)), _p(inds), _p(thresh),
                              C.c_int(len(inds)), C.c_int(num), C.c_int(ht), C.c_int(wd), _p(out))
    return out


def se3_retr(xi, pose):
    out = np.zeros(7, np.float32)
    lib().oracle_se3_retr(_p(np.ascontiguousarray(xi, np.float32)), _p(np.ascontiguousarray(pose, np.float32)), _p(out))
    return out


def se3_rel(pi, pj):
    out = np.zeros(7, np.float32)
    lib().oracle_se3_rel(_p(np.ascontiguousarray(pi, np.float32)), _p(np.ascontiguousarray(pj, np.float32)), _p(out))
    return out


def ref_eigen_solve(A, b, lm, ep):
    """SparseBlock::solve through the reference's vendored Eigen (oracle/_ref, when built)."""
    global _REF
    so = os.path.join(_HERE, "_ref", "libref_eigen_solve.so")
    if not os.path.exists(so):
        return None
    if _REF is None:
        _REF = C.CDLL(so); _REF.ref_eigen_llt_solve.restype = C.c_int
    A = np.ascontiguousarray(A, np.float64); b = np.ascontiguousarray(b, np.float64)
    x = np.zeros_like(b)
    st = _REF.ref_eigen_llt_solve(_p(A), _p(b), C.c_int(len(b)), C.c_float(lm), C.c_float(ep), _p(x))
    return x, st


# ---- projmap / reproject: elementwise fp32 restatements in numpy (test infrastructure only) ---------------------------
def _rel_pose_f32(pi, pj):
    """relSE3 (droid_kernels.cu:97-107): q_ij = q_j (x) conj(q_i), t_ij = t_j - R(q_ij) t_i; quaternion (x,y,z,w)."""
    f = np.float32
    ti, qi, tj, qj = pi[:3].astype(f), pi[3:].astype(f), pj[:3].astype(f), pj[3:].astype(f)
    q = np.array([-qj[3] * qi[0] + qj[0] * qi[3] - qj[1] * qi[2] + qj[2] * qi[1],
                  -qj[3] * qi[1] + qj[1] * qi[3] - qj[2] * qi[0] + qj[0] * qi[2],
                  -qj[3] * qi[2] + qj[2] * qi[3] - qj[0] * qi[1] + qj[1] * qi[0],
                  qj[3] * qi[3] + qj[0] * qi[0] + qj[1] * qi[1] + qj[2] * qi[2]], f)
    return (tj - _rot_f32(q, ti)).astype(f), q


def _rot_f32(q, X):
    """actSO3 (droid_kernels.cu:58-68) on X[...,3]."""
    f = np.float32
    qv = q[:3].astype(f)
    uv = f(2.0) * np.cross(np.broadcast_to(qv, X.shape), X).astype(f)
    return (X + q[3] * uv + np.cross(np.broadcast_to(qv, X.shape), uv)).astype(f)


def projmap(poses, disps, intrinsics, ii, jj):
    """projmap_kernel (droid_kernels.cu:427-516): coords[num,ht,wd,3] (channel 2 zero), valid[num,ht,wd,1]."""
    f = np.float32
    poses, disps = np.asarray(poses, f), np.asarray(disps, f)
    fx, fy, cx, cy = [f(v) for v in intrinsics]
    _, ht, wd = disps.shape
    v, u = np.meshgrid(np.arange(ht, dtype=f), np.arange(wd, dtype=f), indexing="ij")
    coords = np.zeros((len(ii), ht, wd, 3), f); valid = np.zeros((len(ii), ht, wd, 1), f)
    for e, (i, j) in enumerate(zip(ii, jj)):
        t, q = _rel_pose_f32(poses[i], poses[j])
        X = np.stack([(u - cx) / fx, (v - cy) / fy, np.ones_like(u)], -1).astype(f)
        Y = _rot_f32(q, X) + disps[i][..., None] * t
        ok = Y[..., 2] > f(0.01)
        with np.errstate(divide="ignore", invalid="ignore"):
            coords[e, ..., 0] = np.where(ok, fx * (Y[..., 0] / Y[..., 2]) + cx, u)
            coords[e, ..., 1] = np.where(ok, fy * (Y[..., 1] / Y[..., 2]) + cy, v)
        valid[e, ..., 0] = (Y[..., 2] > f(0.25)).astype(f)
    return coords, valid


def reproject(poses, disps, intrinsics, ii, jj, target=None):
    """projective_transform without Jacobians (geom/projective_ops.py:96-125, via depth_video.py:150-158) and the
    motion features of factor_graph.py:203-205.  intrinsics [4] or [B,4].  Returns coords[N,ht,wd,2],
    valid[N,ht,wd,1] (and motn[N,4,ht,wd] when target[N,ht,wd,2] is given)."""
    f = np.float32
    poses, disps = np.asarray(poses, f), np.asarray(disps, f)
    K = np.asarray(intrinsics, f)
    _, ht, wd = disps.shape
    v, u = np.meshgrid(np.arange(ht, dtype=f), np.arange(wd, dtype=f), indexing="ij")
    N = len(ii)
    coords = np.zeros((N, ht, wd, 2), f); valid = np.zeros((N, ht, wd, 1), f)
    for e, (i, j) in enumerate(zip(ii, jj)):
        Ki, Kj = (K, K) if K.ndim == 1 else (K[i], K[j])
        if i == j:
            t, q = np.array([-0.1, 0, 0], f), np.array([0, 0, 0, 1], f)
        else:
            pi, pj = poses[i].copy(), poses[j].copy()
            pi[3:] /= np.sqrt((pi[3:] ** 2).sum(dtype=f)); pj[3:] /= np.sqrt((pj[3:] ** 2).sum(dtype=f))
            t, q = _rel_pose_f32(pi, pj)
        X = np.stack([(u - Ki[2]) / Ki[0], (v - Ki[3]) / Ki[1], np.ones_like(u)], -1).astype(f)
        Y = _rot_f32(q, X) + disps[i][..., None] * t
        Z = np.where(Y[..., 2] < f(0.1), f(1.0), Y[..., 2])
        d = f(1.0) / Z
        coords[e, ..., 0] = Kj[0] * (Y[..., 0] * d) + Kj[2]
        coords[e, ..., 1] = Kj[1] * (Y[..., 1] * d) + Kj[3]
        valid[e, ..., 0] = (Y[..., 2] > f(0.2)).astype(f)
    if target is None:
        return coords, valid
    grid = np.stack([u, v], -1)
    motn = np.concatenate([coords - grid, np.asarray(target, f) - coords], -1).transpose(0, 3, 1, 2)
    return coords, valid, np.clip(motn, -64.0, 64.0).astype(f)
