"""Generates tests/golden/pyref_ba_*.npz by running the REFERENCE's own PyTorch bundle adjustment
(/root/reference/droid_slam/geom/{ba,chol,projective_ops}.py, imported UNMODIFIED) on seeded inputs.

Run in the build container only (needs /root/reference):   python tests/golden/make_pyref_golden.py

The reference modules import `lietorch` and `torch_scatter`, neither of which is installable here
(SURVEY.md §0), so two tiny shims are registered in sys.modules first:
  * lietorch.SE3 — pure-torch restatement of the group ops the BA uses (mul, inv, act on homogeneous
    points, adjT, retr), following thirdparty/lietorch/lietorch/include/se3.h:36-86,134-142 and
    so3.h:29-60,150-165 (quaternion normalised on construction, like SO3's constructor).
  * torch_scatter.scatter_sum — index_add_ along `dim`.
`projective_ops.py:105` hard-codes device="cuda" for the stereo constant; the module-level name
`torch` inside projective_ops is replaced by a proxy whose as_tensor drops the device argument.

The fixtures pin oracle/oracle.cpp (run in fp64 with the documented CUDA/PyTorch divergences switched
off — SURVEY.md §8a table) to the reference's arithmetic: see tests/test_oracle_pyref.py.
Inputs are regenerated from the seed by droid_b200.synth, only outputs are stored.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(REPO, "droid-slam_reserch_b200"))
REF = "/root/reference/droid_slam"


# ----------------------------------------------------------------------------------- shims
def _qmul(a, b):
    ax, ay, az, aw = a.unbind(-1)
    bx, by, bz, bw = b.unbind(-1)
    return torch.stack([aw * bx + ax * bw + ay * bz - az * by,
                        aw * by - ax * bz + ay * bw + az * bx,
                        aw * bz + ax * by - ay * bx + az * bw,
                        aw * bw - ax * bx - ay * by - az * bz], -1)


def _qrot(q, v):
    qv = q[..., :3]
    uv = 2.0 * torch.cross(qv.expand(torch.broadcast_shapes(qv.shape, v.shape)), v.expand(torch.broadcast_shapes(qv.shape, v.shape)), dim=-1)
    qe = q.expand(uv.shape[:-1] + (4,))
    return v + qe[..., 3:4] * uv + torch.cross(qe[..., :3], uv, dim=-1)


class SE3:
    manifold_dim = 6
    embedded_dim = 7

    def __init__(self, data):
        self.data = data

    def _tq(self):
        t, q = self.data[..., :3], self.data[..., 3:]
        return t, q / q.norm(dim=-1, keepdim=True)

    def inv(self):
        t, q = self._tq()
        qi = q * torch.tensor([-1.0, -1.0, -1.0, 1.0], dtype=q.dtype)
        return SE3(torch.cat([-_qrot(qi, t), qi], -1))

    def mul(self, other):
        t1, q1 = self._tq()
        t2, q2 = other._tq()
        q = _qmul(q1, q2)
        q = q / q.norm(dim=-1, keepdim=True)
        return SE3(torch.cat([t1 + _qrot(q1, t2), q], -1))

    def act(self, p):
        t, q = self._tq()
        if p.shape[-1] == 3:
            return _qrot(q, p) + t
        return torch.cat([_qrot(q, p[..., :3]) + t * p[..., 3:4], p[..., 3:4]], -1)

    def __mul__(self, other):
        return self.mul(other) if isinstance(other, SE3) else self.act(other)

    def adjT(self, a):
        t, q = self._tq()
        qi = q * torch.tensor([-1.0, -1.0, -1.0, 1.0], dtype=q.dtype)
        tau, phi = a[..., :3], a[..., 3:]
        shp = torch.broadcast_shapes(t.shape, tau.shape)
        y_tau = _qrot(qi, tau)
        y_phi = _qrot(qi, phi + torch.cross(tau.expand(shp), t.expand(shp), dim=-1))
        return torch.cat([y_tau, y_phi], -1)

    @staticmethod
    def exp(xi):
        tau, phi = xi[..., :3], xi[..., 3:]
        th2 = (phi * phi).sum(-1, keepdim=True)
        th = th2.sqrt()
        small = th < 1e-6  # EPS in lietorch/include/common.h
        ths = torch.where(small, torch.ones_like(th), th)
        imag = torch.where(small, 0.5 - th2 / 48.0 + th2 * th2 / 3840.0, torch.sin(0.5 * ths) / ths)
        real = torch.where(small, 1.0 - th2 / 8.0 + th2 * th2 / 384.0, torch.cos(0.5 * th))
        q = torch.cat([imag * phi, real], -1)
        q = q / q.norm(dim=-1, keepdim=True)
        c1 = torch.where(small, 0.5 - th2 / 24.0, (1.0 - torch.cos(ths)) / (ths * ths))
        c2 = torch.where(small, 1.0 / 6.0 - th2 / 120.0, (ths - torch.sin(ths)) / (ths * ths * ths))
        pt = torch.cross(phi, tau, dim=-1)
        ppt = torch.cross(phi, pt, dim=-1)
        return SE3(torch.cat([tau + c1 * pt + c2 * ppt, q], -1))

    def retr(self, a):
        return SE3.exp(a).mul(self)

    def __getitem__(self, index):
        return SE3(self.data[index])

    def __setitem__(self, index, item):
        self.data[index] = item.data

    @property
    def shape(self):
        return self.data.shape[:-1]


class Sim3(SE3):
    pass


def install_shims():
    lt = types.ModuleType("lietorch")
    lt.SE3, lt.Sim3 = SE3, Sim3
    sys.modules["lietorch"] = lt
    ts = types.ModuleType("torch_scatter")

    def scatter_sum(src, index, dim=-1, dim_size=None):
        shape = list(src.shape)
        shape[dim] = int(dim_size)
        out = torch.zeros(shape, dtype=src.dtype)
        return out.index_add_(dim, index, src)

    ts.scatter_sum = scatter_sum
    sys.modules["torch_scatter"] = ts
    sys.path.insert(0, REF)
    import geom.projective_ops as pops

    class _TorchProxy:
        def __getattr__(self, k):
            return getattr(torch, k)

        @staticmethod
        def as_tensor(data, **kw):
            kw.pop("device", None)
            return torch.as_tensor(data, **kw)

    pops.torch = _TorchProxy()
    import geom.ba as rba
    return rba, pops


# ----------------------------------------------------------------------------------- cases
def cases():
    from droid_b200 import synth
    out = {}
    ii, jj = synth.neighborhood_edges(0, 6, 3)
    out["nbr6"] = synth.make_ba_problem(ii, jj, 6, ht=24, wd=32, seed=101, t0=1)
    rng = np.random.default_rng(7)
    ii, jj = synth.backend_edges(10, 48, rng, rad=2)
    out["back10"] = synth.make_ba_problem(ii, jj, 10, ht=24, wd=32, seed=102, t0=2)
    return out


def run_reference(rba, prob, iters, motion_only=False):
    t = lambda a: torch.from_numpy(np.asarray(a, np.float64))
    t1 = prob.t1
    poses = SE3(t(prob.poses[:t1])[None].clone())
    q = poses.data[..., 3:]
    poses.data[..., 3:] = q / q.norm(dim=-1, keepdim=True)
    disps = t(prob.disps[:t1])[None].clone()
    intr = t(prob.intrinsics)[None, None].repeat(1, t1, 1)
    ii, jj = torch.from_numpy(prob.ii), torch.from_numpy(prob.jj)
    ht, wd = prob.disps.shape[1:]
    # reference layouts: target/weight [B,N,ht,wd,2]; eta [B,M,ht,wd]
    target = t(prob.target).permute(0, 2, 3, 1).contiguous()[None]
    weight = t(prob.weight).permute(0, 2, 3, 1).contiguous()[None]
    eta = t(prob.eta)[None]
    for _ in range(iters):
        if motion_only:
            poses = rba.MoBA(target, weight, eta, poses, disps, intr, ii, jj, fixedp=prob.t0)
        else:
            poses, disps = rba.BA(target, weight, eta, poses, disps, intr, ii, jj, fixedp=prob.t0)
    return poses.data[0].numpy(), disps[0].numpy()


def main():
    rba, _ = install_shims()
    torch.set_default_dtype(torch.float64)  # torch.eye(D) etc. inside chol.py
    for name, prob in cases().items():
        p1, d1 = run_reference(rba, prob, 1)
        p2, d2 = run_reference(rba, prob, 2)
        pm, _ = run_reference(rba, prob, 2, motion_only=True)
        path = os.path.join(HERE, f"pyref_ba_{name}.npz")
        np.savez_compressed(path, poses_it1=p1, disps_it1=d1, poses_it2=p2, disps_it2=d2, poses_moba2=pm)
        print("wrote", path, os.path.getsize(path))


if __name__ == "__main__":
    main()
