"""Runs the UNMODIFIED reference CUDA extension (built for sm_100a by baseline/build_ref_cuda.sh into
baseline/_ref/droid_backends_ref*.so) on a B200 and stores its outputs on seeded inputs as golden vectors.

    gpurun -- python tests/golden/make_ref_cuda_golden.py          # writes gpurun_out/ref_cuda_golden.npz
    cp gpurun_out/ref_cuda_golden.npz tests/golden/                 # committed fixture

Inputs are regenerated from seeds (droid_b200.synth / numpy Generator), only outputs are stored, as fp16
where the op's tolerance allows to keep the fixture small.  tests/test_golden_ref_cuda.py checks both the
CPU oracle (no GPU needed) and the sm_100a kernels (gpu marker) against this file.
"""
import glob
import os
import sys
import time

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(REPO, "droid-slam_reserch_b200"))
sys.path.insert(0, REPO)


def load_ref():
    import importlib.util
    so = glob.glob(os.path.join(REPO, "baseline", "_ref", "droid_backends_ref*.so"))
    if not so:
        raise SystemExit("baseline/_ref/droid_backends_ref*.so missing: run baseline/build_ref_cuda.sh in the build container")
    spec = importlib.util.spec_from_file_location("droid_backends_ref", so[0])
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def lookup_inputs(seed, shape, dtype):
    rng = np.random.default_rng(seed)
    N, h1, w1, h2, w2 = shape
    vol = rng.normal(0, 1, shape).astype(dtype)
    x = rng.uniform(-0.3 * w2, 1.3 * w2, (N, 1, h1, w1))
    y = rng.uniform(-0.3 * h2, 1.3 * h2, (N, 1, h1, w1))
    co = np.concatenate([x, y], 1).astype(np.float32)
    return vol, co


def altcorr_inputs(seed, B=2, S=1, H=12, W=16, C=128):
    rng = np.random.default_rng(seed)
    f1 = rng.normal(0, 1, (B, H, W, C)).astype(np.float32)
    f2 = rng.normal(0, 1, (B, H, W, C)).astype(np.float32)
    co = np.stack([rng.uniform(-4, W + 4, (B, S, H, W)), rng.uniform(-4, H + 4, (B, S, H, W))], -1).astype(np.float32)
    g = rng.normal(0, 1, (B, S, 49, H, W)).astype(np.float32)
    return f1, f2, co, g


def ba_cases():
    from droid_b200 import synth
    ii, jj = synth.neighborhood_edges(0, 8, 3)
    iis, jjs = np.concatenate([ii, np.arange(8)]), np.concatenate([jj, np.arange(8)])
    return {
        "c1": (synth.config_c1(), 1e-4, 0.1, False),
        "c1s": (synth.make_ba_problem(iis, jjs, 8, seed=77, sensor_frac=0.5), 1e-4, 0.1, False),
        "c2": (synth.config_c2(), 1e-4, 0.1, False),
        "c2b": (synth.config_c2(), 1e-5, 1e-2, False),
    }


def moba_case():
    from droid_b200 import synth
    n_kf, n_new = 6, 8
    ii = np.repeat(np.arange(n_kf, n_kf + n_new), 2)
    jj = np.stack([np.arange(n_new) % (n_kf - 1), np.arange(n_new) % (n_kf - 1) + 1], 1).reshape(-1)
    return synth.make_ba_problem(ii, jj, n_kf + n_new, seed=9, t0=n_kf, t1=n_kf + n_new)


def extra_problem():
    """Problem of make_pyref_reproject_golden.py (stereo edges dropped: projmap has no stereo case)."""
    src = open(os.path.join(HERE, "make_pyref_reproject_golden.py")).read()
    ns = {}
    exec(src[src.index("def problem():"):src.index("def main():")], {"np": np}, ns)
    return ns["problem"]()


def main_extra():
    """Second fixture (ref_cuda_golden2.npz): the entry points the first one did not cover — projmap, iproj,
    depth_filter — on the reprojection problem.      gpurun -- python tests/golden/make_ref_cuda_golden.py extra"""
    ref = load_ref()
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    p = extra_problem()
    sel = p.ii != p.jj
    out = {}
    coords, valid = ref.projmap(t(p.poses), t(p.disps), t(p.intrinsics), t(p.ii[sel]), t(p.jj[sel]))
    out["projmap_coords"], out["projmap_valid"] = coords.cpu().numpy(), valid.cpu().numpy()
    n = p.poses.shape[0]
    out["iproj"] = ref.iproj(t(p.poses), t(p.disps), t(p.intrinsics)).cpu().numpy()
    inds = np.arange(n, dtype=np.int64)
    thresh = np.full(n, 0.05, np.float32)
    out["depth_filter"] = ref.depth_filter(t(p.poses), t(p.disps), t(p.intrinsics), t(inds), t(thresh)).cpu().numpy()
    path = os.path.join(REPO, "gpurun_out", "ref_cuda_golden2.npz")
    os.makedirs(os.path.dirname(path), exist_ok=True)
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", {k: v.shape for k, v in out.items()})


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "extra":
        return main_extra()
    ref = load_ref()
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    out = {}
    for name, (p, lm, ep, mo) in ba_cases().items():
        for it in (1, 2):
            poses, disps = t(p.poses), t(p.disps)
            dx, dz = ref.ba(poses, disps, t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta), t(p.ii),
                            t(p.jj), p.t0, p.t1, it, lm, ep, False)
            out[f"ba_{name}_it{it}_dx"] = dx.cpu().numpy()
            out[f"ba_{name}_it{it}_dz"] = dz.cpu().numpy()
            out[f"ba_{name}_it{it}_poses"] = poses.cpu().numpy()[:p.t1]
    p = moba_case()
    poses, disps = t(p.poses), t(p.disps)
    res = ref.ba(poses, disps, t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta), t(p.ii), t(p.jj), p.t0,
                 p.t1, 1, 1e-4, 0.1, True)
    out["moba_it1_dx"] = res[0].cpu().numpy()
    out["moba_it1_poses"] = poses.cpu().numpy()

    from droid_b200 import synth
    p = synth.config_c2()
    n = p.t1
    ii, jj = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    ii, jj = ii.reshape(-1).astype(np.int64), jj.reshape(-1).astype(np.int64)
    disps = p.disps.copy(); disps[3] = 1e3
    for beta in (0.3, 0.25):
        out[f"fd_beta{beta}"] = ref.frame_distance(t(p.poses), t(disps), t(p.intrinsics), t(ii), t(jj), beta).cpu().numpy()

    for tag, shape, dt in (("f16", (2, 12, 16, 12, 16), np.float16), ("f32", (2, 12, 16, 12, 16), np.float32),
                           ("f16b", (1, 24, 32, 6, 8), np.float16)):
        vol, co = lookup_inputs(21, shape, dt)
        out[f"lookup_{tag}"] = ref.corr_index_forward(t(vol), t(co), 3)[0].cpu().numpy()
    vol, co = lookup_inputs(22, (2, 12, 16, 12, 16), np.float32)
    g = np.random.default_rng(23).normal(0, 1, (2, 7, 7, 12, 16)).astype(np.float32)
    out["lookup_bwd_f32"] = ref.corr_index_backward(t(vol), t(co), t(g), 3)[0].cpu().numpy()

    f1, f2, co, g = altcorr_inputs(31)
    out["altcorr_fwd"] = ref.altcorr_forward(t(f1), t(f2), t(co), 3)[0].cpu().numpy()
    g1, g2, gc = ref.altcorr_backward(t(f1), t(f2), t(co), t(g), 3)
    out["altcorr_bwd_f1"], out["altcorr_bwd_f2"] = g1.cpu().numpy(), g2.cpu().numpy()
    out["altcorr_bwd_coords_absmax"] = np.array(float(gc.abs().max()))

    os.makedirs(os.path.join(REPO, "gpurun_out"), exist_ok=True)
    path = os.path.join(REPO, "gpurun_out", "ref_cuda_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", len(out), "arrays")

    # timing of the reference extension on BASELINE config 3 (reported beside our numbers; BASELINE.md B-REF)
    p = synth.config_c3()
    args = [t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta), t(p.ii), t(p.jj)]
    times = []
    for rep in range(3):
        poses, disps = t(p.poses), t(p.disps)
        torch.cuda.synchronize(); t0 = time.time()
        ref.ba(poses, disps, *args, p.t0, p.t1, 2, 1e-5, 1e-2, False)
        torch.cuda.synchronize(); times.append(time.time() - t0)
    print("reference CUDA ext ba c3 (6000 edges, 2 iters): seconds per call", times)
    with open(os.path.join(REPO, "gpurun_out", "ref_cuda_timing.txt"), "w") as f:
        f.write(f"ba_c3_2iters_seconds {times}\n")


if __name__ == "__main__":
    main()
