"""Golden vectors at the configs the headline metric is quoted on (BASELINE.json configs[2], configs[3]).

  python tests/golden/make_c3_golden.py oracle            # HERE (CPU): oracle fp32 + fp64 -> tests/golden/oracle_c3.npz
  gpurun -- python tests/golden/make_c3_golden.py refcuda  # B200: UNMODIFIED reference CUDA extension (baseline/_ref)
                                                           #   -> gpurun_out/ref_cuda_c3.npz  (copy to tests/golden/)

Cases (inputs regenerated from seeds by droid_b200.synth; only outputs are stored):
  c3   512 keyframes / ~6000 edges / 48x64, mono graph              (synth.config_c3())
  c3s  512 keyframes / ~6000 edges / 40x64, 20 % stereo ii==jj edges + sensor-depth prior on 20 % of the pixels
       (SURVEY.md §8d's stereo variant; 40x64 is this fork's EuRoC map size, Notes/Code Logic.md:67-68)
  c4   1536 keyframes / ~20k edges, three fused maps (synth.config_c4()), one iteration
Each case: `ba(..., itrs=k, lm=1e-5, ep=1e-2)` from the SAME initial state for k = 1 and k = 2 (factor_graph.py:297-298);
stored per call: dx[P,6] (full), and a digest of dz[M,HW]: its norm, per-depth-map norms, 4096 sampled entries.
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(REPO, "droid-slam_reserch_b200"))
sys.path.insert(0, REPO)
sys.path.insert(0, HERE)

LM, EP = 1e-5, 1e-2


def cases(which=("c3", "c3s", "c4")):
    from droid_b200 import synth
    out = {}
    if "c3" in which:
        out["c3"] = (synth.config_c3(), (1, 2))
    if "c3s" in which:
        out["c3s"] = (synth.config_c3(seed=1239, stereo_frac=0.2, ht=40, wd=64, sensor_frac=0.2), (1, 2))
    if "c4" in which:
        out["c4"] = (synth.config_c4(), (1,))
    return out


def dz_sample_index(M, HW, n=4096, seed=99):
    rng = np.random.default_rng(seed)
    return rng.integers(0, M, n), rng.integers(0, HW, n)


def digest(prefix, dx, dz, out):
    dz = np.asarray(dz)
    r, c = dz_sample_index(*dz.shape)
    out[f"{prefix}_dx"] = np.asarray(dx, np.float32)
    out[f"{prefix}_dz_norm"] = np.array(np.linalg.norm(dz.astype(np.float64)))
    out[f"{prefix}_dz_rownorm"] = np.linalg.norm(dz.astype(np.float64), axis=1).astype(np.float32)
    out[f"{prefix}_dz_sample"] = dz[r, c].astype(np.float32)


def main_oracle(which):
    from oracle import oracle
    out = {}
    for name, (p, its) in cases(which).items():
        for it in its:
            for tag, dt in (("f32", np.float32), ("f64", np.float64)):
                t = time.time()
                r = oracle.ba(p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, p.jj, p.t0, p.t1,
                              it, LM, EP, dtype=dt)
                print(f"oracle {name} it{it} {tag}: {time.time() - t:.1f}s status {r['status']}", flush=True)
                digest(f"{name}_it{it}_{tag}", r["dx"], r["dz"], out)
    path = os.path.join(HERE, "oracle_c3.npz")
    if os.path.exists(path):                       # cases are generated in separate runs (c4 takes minutes): merge
        old = dict(np.load(path))
        old.update(out)
        out = old
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


def main_refcuda(which):
    import torch
    import make_ref_cuda_golden as G
    ref = G.load_ref()
    import droid_backends
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    rel = lambda a, b: float(np.linalg.norm(np.asarray(a, np.float64) - b) / max(np.linalg.norm(np.asarray(b, np.float64)), 1e-30))
    out, report = {}, []
    opath = os.path.join(HERE, "oracle_c3.npz")
    orc = np.load(opath) if os.path.exists(opath) else {}
    for name, (p, its) in cases(which).items():
        args = [t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta), t(p.ii), t(p.jj)]
        for it in its:
            poses, disps = t(p.poses), t(p.disps)
            torch.cuda.synchronize(); t0 = time.time()
            dx, dz = ref.ba(poses, disps, *args, p.t0, p.t1, it, LM, EP, False)
            torch.cuda.synchronize(); sec = time.time() - t0
            dxr, dzr = dx.cpu().numpy(), dz.cpu().numpy()
            digest(f"{name}_it{it}_ref", dxr, dzr, out)
            out[f"{name}_it{it}_ref_seconds"] = np.array(sec)
            poses, disps = t(p.poses), t(p.disps)
            dxo, dzo = droid_backends.ba(poses, disps, *args, p.t0, p.t1, it, LM, EP, False)
            dxo, dzo = dxo.cpu().numpy(), dzo.cpu().numpy()
            line = dict(case=name, it=it, ref_seconds=round(sec, 2), kernel_vs_ref_dx=rel(dxo, dxr), kernel_vs_ref_dz=rel(dzo, dzr))
            for tag in ("f32", "f64"):
                k = f"{name}_it{it}_{tag}_dx"
                if k in orc:
                    line[f"ref_vs_oracle_{tag}_dx"] = rel(dxr, orc[k])
                    line[f"kernel_vs_oracle_{tag}_dx"] = rel(dxo, orc[k])
                    r, c = dz_sample_index(*dzr.shape)
                    line[f"ref_vs_oracle_{tag}_dz_sample"] = rel(dzr[r, c], orc[f"{name}_it{it}_{tag}_dz_sample"])
                    line[f"kernel_vs_oracle_{tag}_dz_sample"] = rel(dzo[r, c], orc[f"{name}_it{it}_{tag}_dz_sample"])
            print(line, flush=True)
            report.append(line)
    os.makedirs(os.path.join(REPO, "gpurun_out"), exist_ok=True)
    path = os.path.join(REPO, "gpurun_out", "ref_cuda_c3.npz")
    np.savez_compressed(path, **out)
    import json
    with open(os.path.join(REPO, "gpurun_out", "ref_cuda_c3_report.json"), "w") as f:
        json.dump(report, f, indent=1)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    mode = sys.argv[1] if len(sys.argv) > 1 else "oracle"
    which = tuple(sys.argv[2].split(",")) if len(sys.argv) > 2 else ("c3", "c3s", "c4")
    (main_oracle if mode == "oracle" else main_refcuda)(which)
