"""Generates tests/golden/pyref_graph_edges.npz by running the REFERENCE's own edge selection —
/root/reference/droid_slam/factor_graph.py: `add_proximity_factors` (:315-379), `__filter_repeated_edges` (:44-55),
`add_neighborhood_factors` (:302-312), imported UNMODIFIED — on the cases of tests/graph_cases.py.

Run in the build container only (needs /root/reference):   python tests/golden/make_pyref_graph_golden.py

What is stubbed, and why it does not touch the code under test:
  * sys.modules: `lietorch` (pure-torch shim of make_pyref_golden.py), `matplotlib(.pyplot)` (imported at module top,
    unused by these methods; not installed here), `droid_backends` (imported by modules/corr.py at module top; unused);
  * `video`: an object with the attributes the three methods read: `counter.value`, `stereo`, `ht`, `wd`, `disps`, and
    `distance(ii, jj, beta)` = 0.5 (d1 + d2) of the two directions of `frame_distance` exactly as
    depth_video.py:171-181 computes it — with the CPU oracle's `frame_distance`, which tests/test_golden_ref_cuda.py
    pins bit for bit to the reference's CUDA kernel;
  * `FactorGraph.add_factors` (needs the networks' feature buffers) is replaced ON THE INSTANCE by a recorder: the edge
    list handed to it is the output stored here.
`torch.argsort` on CPU is given equal keys (inf) only beyond the point where the loop stops considering them, and equal
finite distances do not occur in these cases (asserted), so the unspecified tie order cannot influence the lists.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(REPO, "droid-slam_reserch_b200"))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))
sys.path.insert(0, HERE)


def load_reference_factor_graph():
    import make_pyref_golden as G
    G.install_shims()                                   # lietorch, torch_scatter, sys.path -> /root/reference/droid_slam
    mpl = types.ModuleType("matplotlib"); plt = types.ModuleType("matplotlib.pyplot"); mpl.pyplot = plt
    sys.modules.setdefault("matplotlib", mpl); sys.modules.setdefault("matplotlib.pyplot", plt)
    sys.modules.setdefault("droid_backends", types.ModuleType("droid_backends"))
    import factor_graph as fg
    assert fg.__file__.startswith("/root/reference/"), fg.__file__
    return fg


class _Counter:
    def __init__(self, v):
        self.value = v


class StubVideo:
    def __init__(self, p, t, stereo, beta_log):
        from oracle import oracle
        self._o, self.p = oracle, p
        self.counter, self.stereo = _Counter(t), stereo
        self.ht, self.wd = 8 * p.disps.shape[1], 8 * p.disps.shape[2]
        self.disps = torch.from_numpy(p.disps)
        self.beta_log = beta_log

    def distance(self, ii, jj, beta=0.3):               # depth_video.py:160-190, bidirectional
        self.beta_log.append(beta)
        ii, jj = ii.reshape(-1).numpy().astype(np.int64), jj.reshape(-1).numpy().astype(np.int64)
        if ii.size == 0:
            return torch.zeros(0)
        poses = np.ascontiguousarray(self.p.poses[:self.counter.value])
        d1 = self._o.frame_distance(poses, self.p.disps, self.p.intrinsics, ii, jj, beta)
        d2 = self._o.frame_distance(poses, self.p.disps, self.p.intrinsics, jj, ii, beta)
        return .5 * (torch.from_numpy(d1) + torch.from_numpy(d2))


def main():
    import graph_cases as GC
    fgm = load_reference_factor_graph()
    out = {}
    for c, case in enumerate(GC.CASES):
        n_kf, t0, t1, rad, nms, beta, thresh, max_factors, stereo, n_old = case
        p, io, jo = GC.case_inputs(case)
        betas = []
        video = StubVideo(p, n_kf, stereo, betas)
        g = fgm.FactorGraph(video, None, device="cpu", max_factors=max_factors)
        # existing edges split over the three lists the method concatenates (:331-332)
        a, b = n_old // 3, 2 * n_old // 3
        T = lambda x: torch.from_numpy(x)
        g.ii, g.jj = T(io[:a]), T(jo[:a])
        g.ii_bad, g.jj_bad = T(io[a:b]), T(jo[a:b])
        g.ii_inac, g.jj_inac = T(io[b:]), T(jo[b:])
        rec = []
        g.add_factors = lambda ii, jj, remove=False: rec.append((ii.clone(), jj.clone()))
        if True:
            try:
                g.add_proximity_factors(t0=t0, t1=t1, rad=rad, nms=nms, beta=beta, thresh=thresh)
                ii, jj = rec[0]
                es = np.stack([ii.numpy(), jj.numpy()], 1).astype(np.int64)
            except Exception as ex:                    # the reference fails on an empty list (unbind of an empty tensor)
                es = np.zeros((0, 2), np.int64)
                out[f"case{c}_reference_raised"] = np.array(type(ex).__name__)
        # distances must not tie among the finite candidates the loop can pick
        gi, gj = [x.reshape(-1) for x in torch.meshgrid(torch.arange(t0, n_kf), torch.arange(t1, n_kf), indexing="ij")]
        d = video.distance(gi, gj, beta=beta).numpy()
        d = d[(gi - rad >= gj).numpy()]                 # the candidates the method keeps (:328)
        fin = np.sort(d[(d <= thresh)])
        assert fin.size < 2 or (np.diff(fin) > 0).all(), "tied candidate distances"
        assert betas[0] == beta
        out[f"case{c}_es"] = es
        print(f"case {c}: {len(es)} edges", out.get(f"case{c}_reference_raised", ""))
    # __filter_repeated_edges (:44-55)
    ii, jj, oi, oj = GC.filter_case()
    p, _, _ = GC.case_inputs(GC.CASES[0])
    g = fgm.FactorGraph(StubVideo(p, 40, False, []), None, device="cpu")
    g.ii, g.jj = torch.from_numpy(oi[:200]), torch.from_numpy(oj[:200])
    g.ii_inac, g.jj_inac = torch.from_numpy(oi[200:]), torch.from_numpy(oj[200:])
    ki, kj = g._FactorGraph__filter_repeated_edges(torch.from_numpy(ii), torch.from_numpy(jj))
    out["filter_ii"], out["filter_jj"] = ki.numpy(), kj.numpy()
    # add_neighborhood_factors (:302-312), mono and stereo
    for stereo in (False, True):
        g = fgm.FactorGraph(StubVideo(p, 40, stereo, []), None, device="cpu")
        rec = []
        g.add_factors = lambda ii, jj, remove=False: rec.append((ii.clone(), jj.clone()))
        g.add_neighborhood_factors(2, 11, r=3)
        out[f"nbr_stereo{int(stereo)}"] = np.stack([rec[0][0].numpy(), rec[0][1].numpy()], 1)
    path = os.path.join(HERE, "pyref_graph_edges.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
