"""Generates tests/golden/pyref_reproject.npz by running the REFERENCE's own `projective_transform`
(/root/reference/droid_slam/geom/projective_ops.py:96-125, imported UNMODIFIED under the lietorch shim of
make_pyref_golden.py) in fp64 on a seeded problem that includes stereo (ii == jj) edges and points behind the
target camera.  Build container only:   python tests/golden/make_pyref_reproject_golden.py
Pins oracle.reproject (tests/test_oracle.py) and, through it, the CUDA kernel (tests/test_gpu_graph.py)."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_pyref_golden as base  # noqa: E402


def problem():
    from droid_b200 import synth
    ii, jj = synth.neighborhood_edges(0, 7, 3)
    ii = np.concatenate([ii, np.arange(7)]); jj = np.concatenate([jj, np.arange(7)])      # + stereo edges
    p = synth.make_ba_problem(ii, jj, 7, ht=24, wd=32, seed=303)
    p.disps[2, :6] = 300.0         # very close points: some land behind / too near the neighbouring cameras
    p.poses[5, 3:] *= 1.001        # a quaternion that drifted off unit norm (the CUDA BA never renormalises)
    return p


def main():
    _, pops = base.install_shims()
    torch.set_default_dtype(torch.float64)      # the stereo constant of projective_ops.py:105 is built from Python floats
    p = problem()
    t = lambda a: torch.from_numpy(np.asarray(a, np.float64))
    poses = base.SE3(t(p.poses)[None])
    intr = t(p.intrinsics)[None, None].repeat(1, p.poses.shape[0], 1)
    ii, jj = torch.from_numpy(p.ii), torch.from_numpy(p.jj)
    coords, valid = pops.projective_transform(poses, t(p.disps)[None], intr, ii, jj)
    path = os.path.join(HERE, "pyref_reproject.npz")
    np.savez_compressed(path, coords=coords[0].numpy().astype(np.float32), valid=valid[0].numpy().astype(np.float32))
    print("wrote", path, os.path.getsize(path), "valid fraction", float(valid.mean()))


if __name__ == "__main__":
    main()
