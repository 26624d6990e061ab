"""Shared by tests/test_gpu_dropin.py and tests/golden/make_dropin_golden.py: loads the reference's OWN
droid_slam/modules/corr.py (unmodified; copied verbatim by __graft_entry__.build() into the git-ignored
baseline/_ref/ref_py/, because /root/reference does not exist on the GPU box) with a chosen `droid_backends`
behind it, and builds the seeded inputs of the drop-in cases.

The cases follow the reference's call sites:
  CorrBlock     factor_graph.py:85,114-116,208 (f16 fmaps under autocast, coords1[1,N,ht,wd,2])
  AltCorrBlock  factor_graph.py:261-277 (update_lowmem: fmaps.view(1, num*rig, ch, ht, wd), indices rig*iis and
                rig*jjs + (iis == jjs) — the stereo right image of frame i is fmaps[i, 1])
"""
import glob
import importlib.util
import os
import sys
import types

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_PY = os.path.join(REPO, "baseline", "_ref", "ref_py", "modules_corr.py")


def load_reference_ext():
    so = glob.glob(os.path.join(REPO, "baseline", "_ref", "droid_backends_ref*.so"))
    if not so:
        return None
    spec = importlib.util.spec_from_file_location("droid_backends_ref", so[0])
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def load_reference_corr(backend, name):
    """Execute the reference's modules/corr.py with `import droid_backends` resolving to `backend`."""
    if not os.path.exists(REF_PY):
        return None
    saved = sys.modules.get("droid_backends")
    sys.modules["droid_backends"] = backend
    try:
        mod = types.ModuleType(name)
        mod.__file__ = REF_PY
        with open(REF_PY) as f:
            exec(compile(f.read(), REF_PY, "exec"), mod.__dict__)
    finally:
        if saved is None:
            sys.modules.pop("droid_backends", None)
        else:
            sys.modules["droid_backends"] = saved
    return mod


def smooth_coords(rng, n, ht, wd, sigma=1.5, shift=6.0, wild=0.02):
    """coords[n,ht,wd,2] (x,y) like reprojections: the pixel grid + a smooth per-edge flow + small noise; a few pixels are
    thrown far outside the image (points behind the camera reproject anywhere)."""
    yy, xx = np.meshgrid(np.arange(ht, dtype=np.float64), np.arange(wd, dtype=np.float64), indexing="ij")
    out = np.empty((n, ht, wd, 2))
    for e in range(n):
        a = rng.normal(0, 0.06, (2, 2)) + np.eye(2)
        t = rng.normal(0, shift, 2)
        out[e, ..., 0] = a[0, 0] * xx + a[0, 1] * yy + t[0]
        out[e, ..., 1] = a[1, 0] * xx + a[1, 1] * yy + t[1]
    out += rng.normal(0, sigma, out.shape)
    far = rng.random((n, ht, wd)) < wild
    out[far] += rng.normal(0, 300.0, (int(far.sum()), 2))
    return out.astype(np.float32)


def corrblock_case(seed=41, n=5, ht=24, wd=32, dtype=np.float16):
    rng = np.random.default_rng(seed)
    f1 = rng.normal(0, 1, (1, n, 128, ht, wd)).astype(dtype)
    f2 = rng.normal(0, 1, (1, n, 128, ht, wd)).astype(dtype)
    co = smooth_coords(rng, n, ht, wd)[None]
    return f1, f2, co


def altcorr_case(seed=43, num=6, rig=2, ht=24, wd=32, dtype=np.float16, S=None):
    """video.fmaps[num, rig, 128, ht, wd]; a chunk of update_lowmem's edges incl. stereo (ii == jj) ones."""
    rng = np.random.default_rng(seed)
    fmaps = rng.normal(0, 1, (num, rig, 128, ht, wd)).astype(dtype)
    ii = np.array([0, 0, 1, 1, 2, 3, 3, 4, 5, 2], np.int64)
    jj = np.array([1, 0, 0, 1, 4, 3, 5, 2, 5, 0], np.int64)
    ii, jj = ii[ii < num], jj[:int((ii < num).sum())]
    co = smooth_coords(rng, len(ii), ht, wd)[None]
    if S is not None:                                   # the 6-D coords form of corr.py:110,128 ([B,N,H,W,S,2])
        co = np.stack([co] + [co + rng.normal(0, 2.0, co.shape).astype(np.float32) for _ in range(S - 1)], -2)
    ii_r = rig * ii
    jj_r = rig * jj + (ii == jj).astype(np.int64) * (rig - 1)
    return fmaps, ii_r, jj_r, co


def sample_index(shape, n=20000, seed=5):
    rng = np.random.default_rng(seed)
    return rng.integers(0, int(np.prod(shape)), n)
