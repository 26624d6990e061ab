#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 hot path (dense BA edge-iterations/s; corr lookups/s beside it).

  python bench.py --gpus N --steps K --warmup W            # this repo's sm_100a path
  python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU path (oracle port) on host cores

Workload (BASELINE.json configs[2], the one `north_star` quotes its target on): stereo-EuRoC-shaped global
backend BA, 512 keyframes, ~6k edges, 48x64 maps, one step = one `droid_backends.ba(..., itrs=2, lm=1e-5,
ep=1e-2)` call exactly as `FactorGraph.update_lowmem` issues it (droid_slam/factor_graph.py:297-298) from
the SAME initial state every step.  value = edges * itrs * steps / time  [edge-iterations/s].
N > 1: edges sharded by source frame over the ranks (one process per GPU), one NCCL all-reduce of the
reduced pose system per iteration (droid_b200/dist.py) — total work fixed => "scaling": "strong".

Prints ONE JSON line on rank 0 (see the task contract): value (inputs resident in HBM), e2e (host buffers,
H2D/D2H inside the timed region), roofline of the dominant kernel (measured live with CUDA events recorded
by the library around it), cpu_baseline (oracle port on a bounded sample), clocks, gpu_launches.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
for _p in (REPO, os.path.join(REPO, "droid-slam_reserch_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

UNIT = "edge-iters/s"
ITRS, LM, EP = 2, 1e-5, 1e-2          # factor_graph.py:297-298
STAGES = ["graph", "edge_prep", "accumulate", "solve", "post", "backsub"]

# BASELINE.json configs[2] (the one the metric is quoted on) and configs[3]; c3m is the round-1 graph (no stereo edges, no
# sensor prior), kept for comparison with BENCH_r01.json
WORKLOADS = {
    "c3": dict(workload="c3: stereo-EuRoC global backend BA (512 keyframes, ~6k edges, 20 % of the frames with a stereo "
                        "ii==jj edge, sensor-depth prior on 20 % of the pixels)", keyframes=512, edges_target=6000, ht=48, wd=64),
    "c3m": dict(workload="c3m: mono global backend BA (512 keyframes, ~6k edges, no stereo edges, no sensor prior; the "
                         "round-1 bench graph)", keyframes=512, edges_target=6000, ht=48, wd=64),
    "c4": dict(workload="c4: multisession BackendAllMaps, 3 fused maps (1536 keyframes, ~20k edges incl. stereo and "
                        "cross-map edges), maps loaded from reconstruction .npy folders by owner rank", keyframes=1536,
               edges_target=20000, ht=48, wd=64),
}


def metric_name(wl):
    w = WORKLOADS[wl]
    return (f"dense BA edge-iterations/s ({'multisession' if wl == 'c4' else 'global'} backend BA, {w['keyframes']} keyframes / "
            f"~{w['edges_target'] // 1000}k edges, itrs=2 per call)")


def workload_config(wl):
    return dict(WORKLOADS[wl], itrs_per_call=ITRS, lm=LM, ep=EP, t0=1)


def make_problem(wl):
    from droid_b200 import synth
    if wl == "c3":
        return synth.config_c3(stereo_frac=0.2, sensor_frac=0.2)
    if wl == "c3m":
        return synth.config_c3()
    return synth.config_c4()


def peaks():
    try:
        with open(os.path.join(REPO, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json, burst copy)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")] + [time.perf_counter()])

    def wait_first(self, timeout=5.0):
        """nvidia-smi needs a few hundred ms before its first row: a short timed region would end before it"""
        t0 = time.perf_counter()
        while self.proc is not None and not self.rows and time.perf_counter() - t0 < timeout:
            time.sleep(0.02)

    def mark(self):
        return time.perf_counter()

    def stop(self, t_begin=None, t_end=None):
        """summary over the samples taken in [t_begin - 0.1 s, t_end + 0.1 s] (all samples when not given)"""
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        rows = self.rows
        if t_begin is not None and t_end is not None:
            sel = [r for r in rows if t_begin - 0.1 <= r[-1] <= t_end + 0.1]
            rows = sel if sel else rows          # a region shorter than the sampling period: fall back to the whole run under load
        for r in rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2])); pw.append(float(r[3]))
            except Exception:
                continue
            for name, col in (("hw_slowdown", 5), ("hw_thermal_slowdown", 6), ("sw_thermal_slowdown", 7), ("sw_power_cap", 8)):
                if len(r) > col and r[col].lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(pw)}


# --------------------------------------------------------------------------------------------- CPU arm
def run_cpu(wl, steps, warmup):
    """The reference's CPU path for this metric = oracle/oracle.cpp (OpenMP restatement of src/droid_kernels.cu ba_cuda,
    fp32 kernels + fp64 solve) on the FULL workload — the same graph, maps and call as the GPU arm (a c3 call takes
    ~10-20 s on 16 cores, a c4 call about a minute)."""
    from oracle import oracle
    p = make_problem(wl)
    cores = oracle.num_threads()
    times = []
    for s in range(warmup + steps):
        t = time.perf_counter()
        oracle.ba(p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, p.jj, p.t0, p.t1, ITRS, LM, EP)
        if s >= warmup:
            times.append(time.perf_counter() - t)
    tot = sum(times)
    sample = f"the full {wl} workload ({p.n_edges} edges, {p.t1} keyframes), {len(times)} timed call(s) of itrs={ITRS}"
    return dict(value=p.n_edges * ITRS * len(times) / tot, unit=UNIT, cores=cores, kind="port", sample=sample,
                ms_per_step=1e3 * tot / len(times), edges=p.n_edges)


def main_reference(args, rank):
    if rank != 0:
        return
    # torchrun exports OMP_NUM_THREADS=1 to every rank; the reference arm is the CPU path "with all the host threads it
    # can use", and only rank 0 runs it — restore the full core count before the OpenMP runtime of the oracle starts
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
    cap = 1 if args.workload == "c4" else 3            # bounded: a c3 call is ~10-20 s of CPU work, a c4 call ~1 min
    steps, warmup = max(1, min(args.steps, cap)), min(args.warmup, 1 if args.workload != "c4" else 0)
    r = run_cpu(args.workload, steps, warmup)
    line = dict(impl="reference", metric=metric_name(args.workload), value=r["value"], unit=UNIT, n_gpus=args.gpus, steps=steps,
                warmup=warmup, ms_per_step=r["ms_per_step"], higher_is_better=True, scaling="strong", vs_baseline=None, dtype="f32",
                data="synthetic", config=dict(workload_config(args.workload), sample=r["sample"], edges=r["edges"]),
                cpu_baseline=dict(value=r["value"], unit=UNIT, cores=r["cores"], kind=r["kind"], sample=r["sample"]),
                e2e=dict(value=r["value"], unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                note="reference's CPU arm = oracle/oracle.cpp (OpenMP restatement of src/droid_kernels.cu ba_cuda, fp32 kernels + "
                     f"fp64 solve) on the full workload; steps capped at {cap} (each call is tens of seconds of CPU work)")
    emit(line)


# --------------------------------------------------------------------------------------------- GPU arm
def _load_ref_ext():
    """The unmodified reference CUDA extension (baseline/_ref, built by baseline/build_ref_cuda.sh) or None."""
    import glob
    import importlib
    refdir = os.path.join(REPO, "baseline", "_ref")
    if not glob.glob(os.path.join(refdir, "droid_backends_ref*.so")):
        return None
    try:
        if refdir not in sys.path:
            sys.path.insert(0, refdir)
        return importlib.import_module("droid_backends_ref")
    except Exception:
        return None


def lookup_bench(dev, steps, warmup, hbm_peak, world=1):
    """Secondary metric: fused 4-level radius-3 CorrBlock lookup on f16 volumes, 96 edges x 48x64 (c2 window).
    One lookup = one (edge, pixel) producing all 4x49 outputs; algorithmic bytes 912 B/lookup."""
    import torch
    import droid_backends
    N, ht, wd = 96, 48, 64
    g = torch.Generator(device=dev); g.manual_seed(7)
    levels = [torch.randn((N, ht, wd, ht >> l, wd >> l), device=dev, dtype=torch.float16, generator=g) for l in range(4)]
    # the CorrBlock path keeps its wide levels in the tiled slice layout (include/droid_b200.h); the nine-name operators and a
    # caller that owns its volumes read row-major ones: both are timed, the CorrBlock path is the block's headline
    mask = droid_backends.corr_tiled_levels(128, ht, wd, 4, torch.float16)
    tiled = [droid_backends.tile_level(v).contiguous() if (mask >> l) & 1 else v for l, v in enumerate(levels)]
    yy, xx = torch.meshgrid(torch.arange(ht, device=dev, dtype=torch.float32), torch.arange(wd, device=dev, dtype=torch.float32),
                            indexing="ij")
    coords = torch.stack([xx, yy], 0)[None].repeat(N, 1, 1, 1) + 6.0 * torch.randn((N, 2, ht, wd), device=dev, generator=g)
    coords = coords.contiguous()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]

    def timed(lv, m):
        for _ in range(warmup):
            droid_backends.corr_pyramid_lookup(lv, coords, 3, tiled_mask=m)
        torch.cuda.synchronize()
        ev[0].record()
        for _ in range(steps):
            droid_backends.corr_pyramid_lookup(lv, coords, 3, tiled_mask=m)
        ev[1].record()
        torch.cuda.synchronize()
        return ev[0].elapsed_time(ev[1]) / steps
    same = bool(torch.equal(droid_backends.corr_pyramid_lookup(levels, coords, 3),
                            droid_backends.corr_pyramid_lookup(tiled, coords, 3, tiled_mask=mask)))
    ms_row = timed(levels, 0)
    ms_loc = timed(tiled, mask)
    ms, lookups = _agg(dev, world, ms_loc, N * ht * wd)      # N > 1: every rank looks up its own 96 frontend edges (replicas)
    gbs = (N * ht * wd) * 912 / (ms_loc * 1e-3) / 1e9       # roofline of THIS rank's kernel
    ref_ms = None
    ref = _load_ref_ext() if world == 1 else None
    if ref is not None:                      # CorrBlock.__call__ as the reference runs it: 4 launches + cat (corr.py:40-50)
        try:
            cl = [(coords / 2 ** l).contiguous() for l in range(4)]
            fn = lambda: torch.cat([ref.corr_index_forward(levels[l], cl[l], 3)[0].view(N, -1, ht, wd) for l in range(4)], 1)
            fn(); torch.cuda.synchronize()
            ev[0].record()
            for _ in range(5):
                fn()
            ev[1].record(); torch.cuda.synchronize()
            ref_ms = ev[0].elapsed_time(ev[1]) / 5
        except Exception:
            ref_ms = None
    return dict(metric="corr lookups/s, frontend path (CorrBlock: fused 4-level r=3 lookup on f16 volumes, 96 edges x 48x64 per GPU; "
                       "volumes 2.4 GB > L2; N > 1 = one such window per rank, no collective)",
                value=lookups / (ms * 1e-3), unit="lookups/s", ms_per_launch=ms, reference_cuda_ms=ref_ms,
                layout=f"CorrBlock's private pyramid: levels {[l for l in range(4) if (mask >> l) & 1]} in 4x8-element tiles",
                ms_per_launch_row_major=ms_row, tiled_equals_row_major=same,
                roofline=dict(bound="hbm", achieved=gbs, peak=hbm_peak, unit="GB/s", frac=gbs / hbm_peak, traffic=None,
                              algorithmic_bytes_per_launch=N * ht * wd * 912))


def _agg(dev, world, ms_local, units_local):
    """(max over ranks of the time, sum over ranks of the units) — edge-parallel blocks have no collective of their own"""
    import torch
    import torch.distributed as dist
    if world == 1:
        return ms_local, units_local
    t = torch.tensor([ms_local, -float(units_local)], device=dev, dtype=torch.float64)
    mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    sm = t.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    return float(mx[0].item()), -float(sm[1].item())


def altcorr_bench(p, dev, rank, world, steps=3, warmup=1):
    """The backend's lookup (FactorGraph.update_lowmem, factor_graph.py:261-290): one step = AltCorrBlock over ALL edges of
    the workload in chunks of 8 source frames, exactly as the reference drives it — `corr_op(coords1[:, v], rig * iis,
    rig * jjs + (iis == jjs))` on the stereo feature buffer fmaps[num, rig = 2, 128, ht, wd] (f16).  Edges are sharded by
    source frame over the ranks (the BA partition); no collective.  One lookup = one (edge, pixel) with its 4 x 49 values."""
    import torch
    from droid_b200 import corr as mirror, dist as bdist, synth
    if os.path.join(REPO, "tests") not in sys.path:
        sys.path.insert(0, os.path.join(REPO, "tests"))
    n, rig = int(p.t1), 2
    _, ht, wd = p.disps.shape
    g = torch.Generator(device=dev); g.manual_seed(33)
    fmaps = torch.randn((n, rig, 128, ht, wd), device=dev, dtype=torch.float16, generator=g)
    bounds = bdist.partition_frames(p.ii, p.jj, n, world)
    sel = bdist.shard_edges(p.ii, bounds, rank)
    ii_l, jj_l = p.ii[sel], p.jj[sel]
    co, _ = synth.reproject(p.poses.astype(np.float64), p.disps.astype(np.float64), p.intrinsics, ii_l, jj_l)
    coords1 = torch.from_numpy(np.nan_to_num(co, nan=0.0, posinf=1e4, neginf=-1e4).astype(np.float32)).to(dev)[None]
    ii_t, jj_t = torch.from_numpy(ii_l).to(dev), torch.from_numpy(jj_l).to(dev)
    chunks = []
    for i in range(0, n, 8):
        v = (ii_t >= i) & (ii_t < i + 8)
        if bool(v.any()):
            chunks.append(v)
    masks = chunks                                        # factor_graph.py:268-272 builds these boolean masks per chunk
    from droid_b200 import graph as bgraph
    chunks = bgraph.edge_chunks(ii_t, n, 8)                   # the same chunks as index lists: no host sync when indexing with them
    assert len(chunks) == len(masks) and all(torch.equal(ii_t[c], ii_t[m]) for c, m in zip(chunks, masks))
    op = mirror.AltCorrBlock(fmaps.view(1, n * rig, 128, ht, wd))

    def step(o, which):
        out = None
        for v in which:
            iis, jjs = ii_t[v], jj_t[v]
            out = o(coords1[:, v], rig * iis, rig * jjs + (iis == jjs).long())
        return out

    for _ in range(warmup):
        step(op, chunks)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        step(op, chunks)
    e1.record()
    torch.cuda.synchronize()
    ms_loc = e0.elapsed_time(e1) / steps
    ms, lookups = _agg(dev, world, ms_loc, len(ii_l) * ht * wd)
    edges = lookups / (ht * wd)
    # the same step (a) with the boolean masks of the unmodified factor_graph.py loop (`self.ii[v]`, `coords1[:, v]`: three host
    # synchronisations per chunk — the device idles while the host waits), (b) the launches alone on prepared chunk arguments
    step(op, masks)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        step(op, masks)
    e1.record()
    torch.cuda.synchronize()
    ms_masks, _ = _agg(dev, world, e0.elapsed_time(e1) / steps, 0)
    import droid_backends
    pre = []
    for ix in chunks:
        iis, jjs = ii_t[ix], jj_t[ix]
        pre.append((coords1[0, ix].contiguous(), (rig * iis).contiguous(), (rig * jjs + (iis == jjs).long()).contiguous()))
    for rep_ in range(2):
        torch.cuda.synchronize()
        e0.record()
        for _ in range(steps if rep_ else 1):
            for c_, a_, b_ in pre:
                droid_backends.altcorr_pyramid_forward(op._hi, a_, b_, c_, 3)
        e1.record()
        torch.cuda.synchronize()
    ms_k, _ = _agg(dev, world, e0.elapsed_time(e1) / steps, 0)
    del pre
    # how many (pixel tile x target cell) GEMMs the kernel needed: a property of the coordinates (one un-timed, counted step)
    droid_backends.COUNT_CELLS = dict(cells=0, tile_levels=0)
    step(op, chunks)
    cnt, droid_backends.COUNT_CELLS = droid_backends.COUNT_CELLS, None
    # the same step on a smooth flow field (affine warp + 1.5 px noise per edge): what co-visible frames of a real map give;
    # the synthetic graph's long-range "loop closures" are not co-visible and scatter the windows of a tile over many cells
    import dropin_cases as D
    coords_w = coords1
    coords1 = torch.from_numpy(D.smooth_coords(np.random.default_rng(5), len(ii_l), ht, wd, wild=0.0)).to(dev)[None]
    step(op, chunks)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        step(op, chunks)
    e1.record()
    torch.cuda.synchronize()
    ms_s, _ = _agg(dev, world, e0.elapsed_time(e1) / steps, 0)
    droid_backends.COUNT_CELLS = dict(cells=0, tile_levels=0)
    step(op, chunks)
    cnt_s, droid_backends.COUNT_CELLS = droid_backends.COUNT_CELLS, None
    coords1 = coords_w
    out = dict(metric="corr lookups/s, backend path (AltCorrBlock over all edges of the workload per step, chunks of 8 source "
                      "frames, 4 levels r=3, f16 stereo feature buffer; chunks given as index lists; edge-sharded, no collective)",
               value=lookups / (ms * 1e-3), unit="lookups/s", ms_per_step=ms, edges=int(edges), chunks_this_rank=len(chunks),
               tflops_algorithmic=edges * 2 * 64 * 128 * 4 * ht * wd / (ms * 1e-3) / 1e12,
               output_gbs=edges * 196 * ht * wd * 4 / (ms * 1e-3) / 1e9,
               ms_per_step_boolean_masks=ms_masks, ms_per_step_launches_only=ms_k,
               how_driven="ms_per_step: the loop of factor_graph.py:261-290 with every chunk given as an index list (graph data, built "
                          "once per graph); ms_per_step_boolean_masks: the unmodified loop (mask indexing = 3 host syncs per chunk, the "
                          "device idles); ms_per_step_launches_only: the fused launches on prepared per-chunk arguments",
               bound="tensor pipe / TMEM drain (2.4 MB written per edge is 2 ms per step at the HBM peak: not HBM bound)",
               cells_per_tile_level=cnt["cells"] / max(cnt["tile_levels"], 1),
               smooth_flow=dict(ms_per_step=ms_s, value=lookups / (ms_s * 1e-3), cells_per_tile_level=cnt_s["cells"] / max(cnt_s["tile_levels"], 1),
                                what="same edges, coordinates = affine warp + 1.5 px noise (co-visible frames)"))
    if world == 1:
        try:                                              # the reference's own classes on the reference extension, same tensors
            sys.path.insert(0, os.path.join(REPO, "tests"))
            import dropin_cases as D
            ref_ext = D.load_reference_ext()
            ref_corr = D.load_reference_corr(ref_ext, "ref_corr_bench") if ref_ext is not None else None
            if ref_corr is not None:
                rop = ref_corr.AltCorrBlock(fmaps.view(1, n * rig, 128, ht, wd))
                sub = chunks[:4]
                step(rop, sub); torch.cuda.synchronize()
                e0.record(); step(rop, sub); e1.record(); torch.cuda.synchronize()
                sub_edges = int(sum(int(v.numel()) for v in sub))
                ref_ms = e0.elapsed_time(e1) * len(ii_l) / sub_edges
                out["reference_cuda_ms_per_step"] = ref_ms
                out["reference_cuda_note"] = (f"unmodified modules/corr.py AltCorrBlock on the reference extension, measured on the first "
                                              f"{len(sub)} chunks ({sub_edges} edges) and scaled to all {len(ii_l)} edges")
        except Exception as ex:
            out["reference_cuda_ms_per_step"] = None
            out["reference_cuda_error"] = f"{type(ex).__name__}: {ex}"[:200]
    del fmaps, op, coords1
    torch.cuda.empty_cache()
    return out


def training_bench(dev, steps=10, warmup=3):
    """Secondary timings at the TartanAir training shape (BASELINE config 5: 4 clips x 22 edges = 88 edges, fp32,
    48x64 maps): CorrBlock path = volume pyramid + 4-level lookup + lookup backward per level; AltCorrBlock path =
    altcorr forward/backward per level on channels-last pyramids (corr.py:91-139)."""
    import torch
    import droid_backends
    N, C_, ht, wd = 88, 128, 48, 64
    g = torch.Generator(device=dev); g.manual_seed(21)
    f1 = torch.randn((N, C_, ht, wd), device=dev, generator=g)
    f2 = torch.randn((N, C_, ht, wd), device=dev, generator=g)
    yy, xx = torch.meshgrid(torch.arange(ht, device=dev, dtype=torch.float32), torch.arange(wd, device=dev, dtype=torch.float32), indexing="ij")
    coords = (torch.stack([xx, yy], 0)[None].repeat(N, 1, 1, 1) + 4.0 * torch.randn((N, 2, ht, wd), device=dev, generator=g)).contiguous()

    def timed(fn):
        for _ in range(warmup):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / steps

    out = {}
    out["corr_volume_f32_ms"] = timed(lambda: droid_backends.corr_volume_pyramid(f1, f2, 4))
    levels = droid_backends.corr_volume_pyramid(f1, f2, 4)
    out["lookup_f32_ms"] = timed(lambda: droid_backends.corr_pyramid_lookup(levels, coords, 3))
    grads = [torch.randn((N, 7, 7, ht, wd), device=dev, generator=g) for _ in range(4)]
    cl = [(coords / 2 ** l).contiguous() for l in range(4)]
    out["lookup_backward_f32_4levels_ms"] = timed(lambda: [droid_backends.corr_index_backward(levels[l], cl[l], grads[l], 3) for l in range(4)])
    del levels
    # AltCorrBlock: channels-last pyramid of fmap2, fmap1 at full resolution
    f1c = f1.permute(0, 2, 3, 1).contiguous()
    pyr = [f2]
    for _ in range(3):
        pyr.append(torch.nn.functional.avg_pool2d(pyr[-1], 2, stride=2))
    pyr = [q.permute(0, 2, 3, 1).contiguous() for q in pyr]
    ca = [(coords / 2 ** l).permute(0, 2, 3, 1)[:, None].contiguous() for l in range(4)]       # [B,S=1,H,W,2]
    out["altcorr_forward_4levels_ms"] = timed(lambda: [droid_backends.altcorr_forward(f1c, pyr[l], ca[l], 3) for l in range(4)])
    ga = [torch.randn((N, 1, 49, ht, wd), device=dev, generator=g) for _ in range(4)]
    out["altcorr_backward_4levels_ms"] = timed(lambda: [droid_backends.altcorr_backward(f1c, pyr[l], ca[l], ga[l], 3) for l in range(4)])
    out["shape"] = "88 edges x 48x64, 128 channels, fp32 (BASELINE config 5)"
    # the unmodified reference CUDA extension on the same tensors (absent .so -> skipped); never on the product path
    try:
        import glob
        import importlib
        refdir = os.path.join(REPO, "baseline", "_ref")
        if glob.glob(os.path.join(refdir, "droid_backends_ref*.so")):
            if refdir not in sys.path:
                sys.path.insert(0, refdir)
            ref = importlib.import_module("droid_backends_ref")
            levels = droid_backends.corr_volume_pyramid(f1, f2, 4)
            r = {}
            r["lookup_f32_ms"] = timed(lambda: [ref.corr_index_forward(levels[l], cl[l], 3) for l in range(4)])
            r["lookup_backward_f32_4levels_ms"] = timed(lambda: [ref.corr_index_backward(levels[l], cl[l], grads[l], 3) for l in range(4)])
            del levels
            r["altcorr_forward_4levels_ms"] = timed(lambda: [ref.altcorr_forward(f1c, pyr[l], ca[l], 3) for l in range(4)])
            r["altcorr_backward_4levels_ms"] = timed(lambda: [ref.altcorr_backward(f1c, pyr[l], ca[l], ga[l], 3) for l in range(4)])
            fm1, fm2 = f1.reshape(1, N, C_, ht * wd), f2.reshape(1, N, C_, ht * wd)

            def ref_volume():                                      # CorrBlock.corr + pyramid as the reference does it (corr.py:30-38,63-71)
                c = torch.matmul(fm1.transpose(2, 3), fm2).view(N, ht, wd, ht, wd) / 4.0
                lv = [c]
                x = c.reshape(N * ht * wd, 1, ht, wd)
                for _ in range(3):
                    x = torch.nn.functional.avg_pool2d(x, 2, stride=2)
                    lv.append(x)
                return lv
            r["corr_volume_f32_ms"] = timed(ref_volume)
            out["reference_cuda_same_shapes"] = r
    except Exception as ex:
        out["reference_cuda_same_shapes"] = dict(unavailable=f"{type(ex).__name__}: {ex}"[:200])
    return out


def graph_bench(p, dev, steps=5):
    """Secondary metrics on the c3 keyframe buffer: `frame_distance` over all 512^2 ordered pairs (one direction;
    DepthVideo.distance calls it twice) and the whole GPU edge selection of add_proximity_factors (both directions of
    frame_distance + masking + NMS + sort + greedy pick, factor_graph.py:315-379)."""
    import torch
    import droid_backends
    from droid_b200 import graph
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    poses, disps, intr = t(p.poses), t(p.disps), t(p.intrinsics)
    n = p.t1
    ii, jj = torch.meshgrid(torch.arange(n), torch.arange(n), indexing="ij")
    ii, jj = ii.reshape(-1).to(dev), jj.reshape(-1).to(dev)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    droid_backends.frame_distance(poses, disps, intr, ii, jj, 0.25)
    graph.proximity_edges(poses, disps, intr, n, 0, 0, 2, 2, 0.25, 16.0, 16 * n, False)
    torch.cuda.synchronize()
    ev[0].record()
    for _ in range(steps):
        droid_backends.frame_distance(poses, disps, intr, ii, jj, 0.25)
    ev[1].record()
    for _ in range(steps):
        gi, _ = graph.proximity_edges(poses, disps, intr, n, 0, 0, 2, 2, 0.25, 16.0, 16 * n, False)
    ev[2].record()
    torch.cuda.synchronize()
    fd_ms, px_ms = ev[0].elapsed_time(ev[1]) / steps, ev[1].elapsed_time(ev[2]) / steps
    ref_ms = None
    ref = _load_ref_ext()
    if ref is not None:
        try:
            ref.frame_distance(poses, disps, intr, ii, jj, 0.25); torch.cuda.synchronize()
            ev[0].record()
            for _ in range(steps):
                ref.frame_distance(poses, disps, intr, ii, jj, 0.25)
            ev[1].record(); torch.cuda.synchronize()
            ref_ms = ev[0].elapsed_time(ev[1]) / steps
        except Exception:
            ref_ms = None
    return dict(frame_distance=dict(pairs=int(ii.numel()), ms_per_call=fd_ms, value=ii.numel() / (fd_ms * 1e-3), unit="pairs/s",
                                    reference_cuda_ms=ref_ms),
                proximity_edges=dict(keyframes=int(n), ms_per_call=px_ms, edges_selected=int(gi.numel()),
                                     what="add_proximity_factors edge list on the GPU (rad=2, nms=2, beta=0.25, thresh=16)"))


def reference_cuda_leg(p, dev, steps=3):
    """Extra reported baseline (north_star: "the reference CUDA extension is also reported where it builds for
    sm_100"): the UNMODIFIED reference extension, built by baseline/build_ref_cuda.sh as `droid_backends_ref`, on the
    same c3 problem, same call, CUDA-event timed.  Absent .so -> None.  Never on the product path."""
    import glob
    import importlib
    import torch
    refdir = os.path.join(REPO, "baseline", "_ref")
    if not glob.glob(os.path.join(refdir, "droid_backends_ref*.so")):
        return None
    try:
        if refdir not in sys.path:
            sys.path.insert(0, refdir)
        ref = importlib.import_module("droid_backends_ref")
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        poses0, disps0 = t(p.poses), t(p.disps)
        intr, sens, target, weight, eta, ii, jj = t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta), t(p.ii), t(p.jj)
        poses, disps = poses0.clone(), disps0.clone()

        def step():
            poses.copy_(poses0); disps.copy_(disps0)
            ref.ba(poses, disps, intr, sens, target, weight, eta, ii, jj, p.t0, p.t1, ITRS, LM, EP, False)
        step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        return dict(value=p.n_edges * ITRS / (ms * 1e-3), unit=UNIT, ms_per_step=ms, steps=steps,
                    what="unmodified reference src/droid_kernels.cu ba_cuda rebuilt for sm_100a (host Eigen solve inside), same B200")
    except Exception as ex:  # the reference extension is a reported extra, never a dependency
        return dict(unavailable=f"{type(ex).__name__}: {ex}"[:200])


def frontend_bench(dev, steps=50):
    """BASELINE configs[1]: the local BA of the frontend window (synth.config_c2: 25 keyframes, ~100 edges, 48x64,
    2 iterations per call, `factor_graph.py:199-244`).  Latency-bound: 7 launches per iteration and a one-CTA solve.
    The reference's CUDA op (host Eigen solve + blocking copies inside) is timed on the same tensors when it is built."""
    import glob
    import importlib
    import torch
    import droid_backends
    from droid_b200 import synth
    p = synth.config_c2()
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    poses0, disps0 = t(p.poses), t(p.disps)
    a = [t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta), t(p.ii), t(p.jj)]
    poses, disps = poses0.clone(), disps0.clone()

    def timed(fn, n):
        fn(); fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    def ours():
        poses.copy_(poses0); disps.copy_(disps0)
        droid_backends.ba(poses, disps, *a, p.t0, p.t1, 2, 1e-4, 0.1, False)
    ms = timed(ours, steps)
    out = dict(workload="c2: frontend window BA", keyframes=int(p.t1), edges=int(p.n_edges), itrs_per_call=2, ms_per_call=ms,
               value=p.n_edges * 2 / (ms * 1e-3), unit=UNIT)
    refdir = os.path.join(REPO, "baseline", "_ref")
    if glob.glob(os.path.join(refdir, "droid_backends_ref*.so")):
        try:
            if refdir not in sys.path:
                sys.path.insert(0, refdir)
            ref = importlib.import_module("droid_backends_ref")

            def theirs():
                poses.copy_(poses0); disps.copy_(disps0)
                ref.ba(poses, disps, *a, p.t0, p.t1, 2, 1e-4, 0.1, False)
            out["reference_cuda_ms"] = timed(theirs, 10)
        except Exception as ex:
            out["reference_cuda_ms"] = None
            out["reference_cuda_error"] = f"{type(ex).__name__}: {ex}"[:200]
    return out


def volume_bench(dev, steps, warmup, hbm_peak):
    """Secondary metric: CorrBlock construction — all-pairs 128-d correlation + 4-level pyramid, f16, 48 edges x 48x64
    (corr.py:30-38,63-71).  Output-write bound: algorithmic bytes = pyramid written once (25.1 MB/edge) + fmaps read."""
    import torch
    import droid_backends
    N, C_, ht, wd = 48, 128, 48, 64
    g = torch.Generator(device=dev); g.manual_seed(11)
    f1 = torch.randn((N, C_, ht, wd), device=dev, dtype=torch.float16, generator=g)
    f2 = torch.randn((N, C_, ht, wd), device=dev, dtype=torch.float16, generator=g)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]

    def timed(tiled):
        for _ in range(warmup):
            droid_backends.corr_volume_pyramid(f1, f2, 4, tiled=tiled)
        torch.cuda.synchronize()
        ev[0].record()
        for _ in range(steps):
            droid_backends.corr_volume_pyramid(f1, f2, 4, tiled=tiled)
        ev[1].record()
        torch.cuda.synchronize()
        return ev[0].elapsed_time(ev[1]) / steps
    ms_row = timed(False)                    # the reference's row-major volumes (one GEMM launch per level)
    ms = timed(True)                         # CorrBlock's private pyramid: tiled levels 0-1, all levels from one launch
    hw = ht * wd
    out_bytes = N * hw * sum((ht >> l) * (wd >> l) for l in range(4)) * 2
    in_bytes = 2 * N * C_ * hw * 2
    flops = 2.0 * N * hw * C_ * hw           # the level-0 product (the pooled levels are means of its accumulators)
    gbs = (out_bytes + in_bytes) / (ms * 1e-3) / 1e9

    def ref_build():                         # CorrBlock.__init__ of the reference in f16 (corr.py:23-38,63-71): cuBLAS + 3 avg_pool2d
        a = (f1.reshape(N, C_, hw) / 4.0).transpose(1, 2)
        c = torch.matmul(a, f2.reshape(N, C_, hw) / 4.0).reshape(N * hw, 1, ht, wd)
        lv = [c]
        for _ in range(3):
            c = torch.nn.functional.avg_pool2d(c, 2, stride=2)
            lv.append(c)
        return lv
    ref_ms = None
    try:
        ref_build(); torch.cuda.synchronize()
        ev[0].record()
        for _ in range(5):
            ref_build()
        ev[1].record(); torch.cuda.synchronize()
        ref_ms = ev[0].elapsed_time(ev[1]) / 5
    except Exception:
        pass
    return dict(metric="CorrBlock build edges/s (all-pairs 128-d GEMM + 4-level pyramid, f16, 48 edges x 48x64)",
                value=N / (ms * 1e-3), unit="edges/s", ms_per_call=ms, tflops=flops / (ms * 1e-3) / 1e12,
                layout="CorrBlock's private pyramid (levels 0-1 in 4x8-element tiles, pooled levels from the level-0 launch)",
                ms_per_call_row_major=ms_row,
                reference_torch_ms=ref_ms,
                roofline=dict(bound="hbm", achieved=gbs, peak=hbm_peak, unit="GB/s", frac=gbs / hbm_peak, traffic=None,
                              algorithmic_bytes_per_call=out_bytes + in_bytes))


def main_gpu(args, rank, world):
    import torch
    import torch.distributed as dist
    import droid_backends
    from droid_b200 import _capi, synth

    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _capi.lib()
    hbm_peak, peak_src = peaks()

    wl = args.workload
    p = make_problem(wl)
    N, HW, M, P = p.n_edges, p.disps.shape[1] * p.disps.shape[2], p.eta.shape[0], p.t1 - p.t0
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    maps_note = None
    if wl == "c4":
        # BackendAllMaps.py:43-61,101-138: the three maps come from reconstruction folders on disk.  Rank 0 writes them in the
        # reference's format (droid.py:92-106), every rank loads the replicated arrays whole and the disparity maps of the
        # keyframes it OWNS only (droid_b200/maps.py); rows it does not own stay zero and are never read by its kernels.
        import tempfile
        from droid_b200 import dist as bdist, maps
        root = os.path.join(tempfile.gettempdir(), "db200_bench_c4_maps")
        kf = p.t1 // 3
        if rank == 0:
            for m in range(3):
                sl = slice(m * kf, (m + 1) * kf)
                maps.save_reconstruction(os.path.join(root, f"map{m + 1}"), tstamps=np.arange(kf, dtype=np.float64), poses=p.poses[sl],
                                         disps=p.disps[sl], intrinsics=np.tile(p.intrinsics, (kf, 1)))
        if world > 1:
            dist.barrier()
        lo, hi = (0, p.t1) if world == 1 else maps.owned_range_for_rank(p.ii, p.jj, p.disps.shape[0], rank, world)
        got = maps.load_maps(root, every=1, own_lo=lo, own_hi=hi, device=dev, names=("tstamps", "poses", "intrinsics", "disps"))
        poses0 = got["poses"].float().contiguous()
        disps0 = torch.zeros(p.disps.shape, dtype=torch.float32, device=dev)
        disps0[lo:hi] = got["disps"]
        if world > 1:                                   # the benchmark resets disps from disps0 every step: assemble it once
            dist.all_reduce(disps0)
        assert torch.equal(poses0.cpu(), torch.from_numpy(p.poses)) and torch.equal(disps0.cpu(), torch.from_numpy(p.disps))
        maps_note = f"maps loaded from {root} (3 folders, .npy, heavy arrays by owner: rank {rank} frames [{lo},{hi}))"
    else:
        poses0, disps0 = t(p.poses), t(p.disps)
    intr, sens = t(p.intrinsics), t(p.disps_sens)

    if world == 1:
        target, weight, eta, ii, jj = t(p.target), t(p.weight), t(p.eta), t(p.ii), t(p.jj)
        poses, disps = poses0.clone(), disps0.clone()

        def step():
            poses.copy_(poses0); disps.copy_(disps0)
            return droid_backends.ba(poses, disps, intr, sens, target, weight, eta, ii, jj, p.t0, p.t1, ITRS, LM, EP, False)
        runner = None
    else:
        from droid_b200 import dist as bdist
        runner = bdist.ShardedBA(p.ii, p.jj, p.t0, p.t1, rank, world, dev)
        sel = runner.local_edges
        target, weight = t(p.target[sel]), t(p.weight[sel])
        eta = t(p.eta)
        poses, disps = poses0.clone(), disps0.clone()

        def step():
            poses.copy_(poses0); disps.copy_(disps0)
            return runner.ba(poses, disps, intr, sens, target, weight, eta, ITRS, LM, EP)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    parity = None
    if world > 1:
        dx_sh, dz_sh = step()
        dx_sh = dx_sh.clone()
        if rank == 0:                                   # one un-timed single-GPU call on the full graph for comparison
            po1, di1 = poses0.clone(), disps0.clone()
            tg1, wg1 = t(p.target), t(p.weight)
            dx1, dz1 = droid_backends.ba(po1, di1, intr, sens, tg1, wg1, eta, t(p.ii), t(p.jj), p.t0, p.t1, ITRS, LM, EP, False)
            rel = lambda a, b: float((a.double() - b.double()).norm() / b.double().norm())
            parity = dict(dx_rel_vs_single=rel(dx_sh, dx1), dz_rel_vs_single=rel(dz_sh, dz1),
                          poses_maxabs_vs_single=float((poses - po1).abs().max()), disps_rel_vs_single=rel(disps, di1),
                          note="edge-sharded call against droid_backends.ba on one GPU, same inputs; the difference is the fp32 "
                               "regrouping of the per-frame sums (test tolerance 3e-4, tests/test_gpu_dist.py)")
            del tg1, wg1, po1, di1
            torch.cuda.empty_cache()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    if rank == 0:
        sampler.wait_first()

    # ---- timed region: device-resident inputs
    t_clk0 = sampler.mark()
    lib.db200_profile_enable(1)
    l0 = lib.db200_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    launches = lib.db200_launch_count() - l0
    st_ms = (C.c_double * len(STAGES))(); st_n = (C.c_int * len(STAGES))()
    lib.db200_profile_collect(st_ms, st_n)
    lib.db200_profile_enable(0)
    t_clk1 = sampler.mark()
    if world > 1:
        tt = torch.tensor([ms_total], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_total = float(tt.item())
    ms_step = ms_total / args.steps
    value = N * ITRS * args.steps / (ms_total * 1e-3)

    # ---- e2e: host (pinned) buffers in, results out, every step, through the same public call
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    if world == 1:
        h_in = [pin(a) for a in (p.target, p.weight, p.eta, p.ii, p.jj, p.poses, p.disps)]
    else:
        # every rank uploads ITS shard: target/weight of its edges, and the eta / disparity rows of the frames it owns (the
        # only rows its kernels read — SURVEY.md §8e: "disps/targets/weights/eta sharded with their owner"); poses replicated
        f_lo, f_hi = runner.own_lo, (p.disps.shape[0] if rank == world - 1 else runner.own_hi)
        e_lo, e_hi = int(np.searchsorted(runner.kx_all, f_lo)), int(np.searchsorted(runner.kx_all, f_hi))
        eta_full, disps_full = t(p.eta), t(p.disps)
        h_in = [pin(a) for a in (p.target[sel], p.weight[sel], p.eta[e_lo:e_hi], p.ii[sel], p.jj[sel], p.poses, p.disps[f_lo:f_hi])]
    # results read back every step: poses and dx (replicated) and the disparity rows this rank owns (all of them at N = 1)
    h_poses, h_disps = torch.empty_like(h_in[5]), torch.empty(h_in[6].shape, dtype=torch.float32).pin_memory()
    h_dx = torch.empty((P, 6), dtype=torch.float32).pin_memory()
    h2d = sum(a.numel() * a.element_size() for a in h_in)
    d2h = sum(a.numel() * a.element_size() for a in (h_poses, h_disps, h_dx))
    if world > 1:                                      # whole-job bytes per step: summed over the ranks
        bb = torch.tensor([h2d, d2h], device=dev, dtype=torch.int64)
        dist.all_reduce(bb, op=dist.ReduceOp.SUM)
        h2d, d2h = int(bb[0].item()), int(bb[1].item())
    # Every step uploads ITS inputs from pinned host memory and reads ITS result back; the upload of step k+1 runs on
    # a copy stream while step k's kernels run (droid_b200/staging.py) — the caller still synchronises on every result.
    from droid_b200.staging import HostStagedCalls
    staged = HostStagedCalls(dev, h_in, depth=2)

    def e2e_step(upload_next):
        if upload_next:
            staged.upload(h_in)
        tg, wg, et, i_, j_, po, di = staged.acquire()
        if world == 1:
            dx, _ = droid_backends.ba(po, di, intr, sens, tg, wg, et, i_, j_, p.t0, p.t1, ITRS, LM, EP, False)
        else:
            eta_full[e_lo:e_hi].copy_(et); disps_full[f_lo:f_hi].copy_(di)      # device-to-device, into the full-size tensors
            di = disps_full
            dx, _ = runner.ba(po, di, intr, sens, tg, wg, eta_full, ITRS, LM, EP)
        h_poses.copy_(po, non_blocking=True); h_dx.copy_(dx, non_blocking=True)
        h_disps.copy_(di if world == 1 else di[f_lo:f_hi], non_blocking=True)
        staged.release()
        torch.cuda.current_stream().synchronize()      # the caller reads the result every step

    e2e_steps = max(3, min(args.steps, 10))
    staged.upload(h_in)
    for _ in range(2):
        e2e_step(True)
    barrier()                                          # the pre-loaded upload of timed step 0 is complete here ...
    e0.record()
    for k in range(e2e_steps):
        e2e_step(True)                                 # ... and every timed step issues exactly one upload
    torch.cuda.current_stream().wait_stream(staged.copy_stream)   # the K-th upload must finish inside the region too:
    e1.record()                                        # K uploads, K calls, K read-backs between e0 and e1
    barrier()
    ms_e2e = e0.elapsed_time(e1)
    if world > 1:
        tt = torch.tensor([ms_e2e], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_e2e = float(tt.item())
    e2e_value = N * ITRS * e2e_steps / (ms_e2e * 1e-3)
    # sampled from before the warm-up to here (nvidia-smi, 100 ms period); the summary prefers the samples inside the
    # device-resident timed region and falls back to the whole loaded interval when that region is shorter than a period
    clocks = sampler.stop(t_clk0, t_clk1) if rank == 0 else None

    # ---- edge-parallel secondary blocks: every rank works on its share (no collective inside the timed loops)
    n_loc = int(target.shape[0])
    del staged, h_in
    if not args.ba_only:
        del target, weight
        torch.cuda.empty_cache()
    blocks = {}
    if not args.ba_only:
        blocks["altcorr"] = altcorr_bench(p, dev, rank, world)
        if wl != "c4":
            blocks["lookup"] = lookup_bench(dev, 20, 3, hbm_peak, world)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant HBM-bound kernel (accumulate): algorithmic bytes per launch / measured launch time
    m_loc = M if world == 1 else runner.n_local_frames
    accum_bytes = 16 * HW * n_loc + 20 * HW * m_loc          # target+weight per edge-pixel; disp,eta,sens in + Q,w out per frame-pixel
    stage = {STAGES[i]: dict(ms_per_iter=(st_ms[i] / st_n[i]) if st_n[i] else None, intervals=int(st_n[i])) for i in range(len(STAGES))}
    acc_ms = stage["accumulate"]["ms_per_iter"]
    achieved = accum_bytes / (acc_ms * 1e-3) / 1e9 if acc_ms else None
    iter_bytes = 20 * HW * N + 16 * HW * M + 56 * P       # SURVEY.md §8d whole-iteration figure (block-sparse system term omitted)
    iter_ms = ms_step / ITRS
    roofline = dict(bound="hbm", kernel="ba_schur_kernel (accumulate stage: reprojection/Jacobians + tensor-core Schur products; 4 degree-bucket launches)", achieved=achieved, peak=hbm_peak, unit="GB/s",
                    frac=(achieved / hbm_peak) if achieved else None, traffic=None, peak_source=peak_src,
                    algorithmic_bytes_per_launch=accum_bytes, ms_per_launch=acc_ms,
                    whole_iteration=dict(algorithmic_bytes=iter_bytes, ms=iter_ms,
                                         achieved=iter_bytes / (iter_ms * 1e-3) / 1e9,
                                         frac=iter_bytes / (iter_ms * 1e-3) / 1e9 / hbm_peak))
    # the solve is FP64-pipe / latency bound, not HBM bound: reported against the measured FP64 rate of this part
    # (DMMA m8n8k4 and register-only DFMA both reach 64 FMA/clk/SM, tools/mb_tile.cu: x 148 SMs x 1.965 GHz x 2 = 37.2 TFLOP/s)
    sol_ms = stage["solve"]["ms_per_iter"]
    n_sys = 6 * P
    sol_flops = n_sys ** 3 / 3.0 + 2.0 * n_sys ** 2
    fp64_peak = float(lib.db200_debug_fp64_peak_gflops(C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    if not fp64_peak > 0:
        fp64_peak = None
    roofline["solver"] = dict(kernel="chol_solve_kernel (cooperative tiled fp64 Cholesky, trailing updates on DMMA + both substitutions)", bound="fp64 pipe / panel-chain latency",
                              gflops=(sol_flops / (sol_ms * 1e-3) / 1e9) if sol_ms else None, peak_gflops=fp64_peak,
                              peak_source="measured in this run: register-only mma.sync.m8n8k4.f64 (DMMA), db200_debug_fp64_peak_gflops",
                              frac=(sol_flops / (sol_ms * 1e-3) / 1e9 / fp64_peak) if (sol_ms and fp64_peak) else None, ms_per_launch=sol_ms,
                              share_of_iteration=(sol_ms / iter_ms) if sol_ms else None)
    traffic_file = os.path.join(REPO, "profiles", "traffic.json")
    if os.path.exists(traffic_file):
        try:
            roofline["traffic"] = json.load(open(traffic_file)).get("ba_schur_kernel_bytes_per_iteration")
        except Exception:
            pass

    line = dict(metric=metric_name(wl), value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=max(args.warmup, 3),
                ms_per_step=ms_step, higher_is_better=True, scaling="strong", vs_baseline=None, dtype="f32 (fp64 reduced system + solve)",
                data="synthetic",
                config=dict(workload_config(wl), edges=N, depth_maps=M, poses=P, parallelism=f"edge-shard x{world}" if world > 1 else "single",
                            stereo_edges=int((p.ii == p.jj).sum()), sensor_prior_pixels=float((p.disps_sens > 0).mean()),
                            timed_step="poses/disps reset from device copies (2 small D2D copies) + one droid_backends.ba call",
                            l2=f"inputs ({16 * HW * N / 1e6:.0f} MB of target/weight per iteration) larger than the 126 MB L2; no explicit flush"),
                e2e=dict(value=e2e_value, unit=UNIT, h2d_bytes_per_step=h2d, d2h_bytes_per_step=d2h, steps=e2e_steps,
                         ms_per_step=ms_e2e / e2e_steps,
                         staging="double-buffered: upload of step k+1 (copy stream) overlaps the kernels of step k; "
                                 "result of every step read back and synchronised; at N > 1 every rank uploads its shard and reads "
                                 "back poses, dx and the disparity rows it owns, bytes summed over the ranks"),
                gpu_launches=int(launches), roofline=roofline, stages=stage, clocks=clocks)
    if parity is not None:
        line["parity"] = parity
    if maps_note:
        line["config"]["maps"] = maps_note
    line.update(blocks)
    if world == 1 and not args.ba_only and wl != "c4":
        line["corr_volume"] = volume_bench(dev, 10, 3, hbm_peak)
        line["graph"] = graph_bench(p, dev)
        line["training_shapes"] = training_bench(dev)
        line["frontend"] = frontend_bench(dev)
        try:
            tr = json.load(open(traffic_file)).get("corr_lookup_r3_direct_kernel_bytes_per_launch")
            line["lookup"]["roofline"]["traffic"] = tr
            # what the kernel actually moves (ncu dram bytes) over the live launch time: the 16-byte window rows cost
            # whole 64-byte DRAM bursts, so the DRAM interface is far busier than the algorithmic figure says
            line["lookup"]["roofline"]["dram_gbs_actual_traffic"] = tr / (line["lookup"]["ms_per_launch"] * 1e-3) / 1e9
            line["lookup"]["roofline"]["dram_frac_actual_traffic"] = line["lookup"]["roofline"]["dram_gbs_actual_traffic"] / hbm_peak
        except Exception:
            pass
    if world == 1 and not args.no_cpu:
        rc = reference_cuda_leg(p, dev, steps=1 if wl == "c4" else 3)
        if rc is not None:
            line["reference_cuda"] = rc
        r = run_cpu(wl, 1, 0)
        line["cpu_baseline"] = dict(value=r["value"], unit=UNIT, cores=r["cores"], kind=r["kind"], sample=r["sample"])
    emit(line)
    if world > 1:
        dist.destroy_process_group()


_RESULT_FD = None


def _claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner to stdout when
    NCCL_DEBUG=VERSION is exported on the box): keep a private copy of the real stdout for the result line and point
    fd 1 at stderr for everything else."""
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    text = json.dumps(line) + "\n"
    if _RESULT_FD is None:
        sys.stdout.write(text); sys.stdout.flush()
    else:
        os.write(_RESULT_FD, text.encode())


def main():
    _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS), help="c3 = BASELINE configs[2] (default), c4 = configs[3]")
    ap.add_argument("--ba-only", action="store_true", help="skip the secondary blocks (lookups, volume build, graph, training shapes)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if args.impl == "reference":
        return main_reference(args, rank)
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if args.gpus > 1 and world == 1:
        raise SystemExit("launch multi-GPU runs with torch.distributed.run (one process per GPU)")
    main_gpu(args, rank, world)


if __name__ == "__main__":
    main()
