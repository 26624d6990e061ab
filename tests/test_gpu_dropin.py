"""Drop-in tests THROUGH THE REFERENCE'S OWN PYTHON: droid_slam/modules/corr.py is executed unmodified twice on the
same B200 — once with `droid_backends` = this repo's shim (sm_100a kernels behind the C-ABI), once with the UNMODIFIED
reference CUDA extension rebuilt for sm_100a (baseline/_ref) — and the results are compared.  The repo's own mirror
(droid_b200.corr, fused build/lookup entry points) is checked against the same reference run.

Tolerances (north_star): lookup values and indices bit-exact; altcorr / fp32 correlation values 1e-4 of the value scale;
f16 volumes built by the fused GEMM: see test_mirror_corrblock_f16 (pooled levels are rounded once, not per level).
"""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import dropin_cases as D  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    import droid_backends
    ref_ext = D.load_reference_ext()
    if ref_ext is None or not os.path.exists(D.REF_PY):
        pytest.skip("baseline/_ref (reference extension + its modules/corr.py) not installed: run __graft_entry__.build() "
                    "where /root/reference exists")
    ours = D.load_reference_corr(droid_backends, "ref_corr_on_b200")
    theirs = D.load_reference_corr(ref_ext, "ref_corr_on_ref_ext")
    return ours, theirs


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _bits(t):
    return t.contiguous().view(torch.int16 if t.dtype == torch.float16 else torch.int32)


def test_reference_corrblock_f16_bit_exact(mods):
    """CorrBlock as the frontend builds it (f16 fmaps): torch builds the pyramid in both runs, every lookup goes through
    droid_backends.corr_index_forward -> bit-exact, all 4 levels, windows crossing the border and far outside."""
    ours, theirs = mods
    f1, f2, co = D.corrblock_case(ht=48, wd=64, n=3)
    a = ours.CorrBlock(_t(f1), _t(f2))(_t(co))
    b = theirs.CorrBlock(_t(f1), _t(f2))(_t(co))
    assert a.shape == b.shape == (1, 3, 196, 48, 64) and a.dtype == b.dtype == torch.float16
    assert torch.equal(_bits(a), _bits(b))


def test_reference_corrblock_cat_getitem_and_backward_fp32(mods):
    """CorrBlock.cat / __getitem__ (factor_graph.py:116,153) and CorrSampler.backward through autograd down to the fmaps."""
    ours, theirs = mods
    f1, f2, co = D.corrblock_case(seed=47, dtype=np.float32, n=4, ht=16, wd=24)
    w = _t(np.random.default_rng(1).normal(0, 1, (1, 4, 196, 16, 24)).astype(np.float32))
    res = []
    for m in (ours, theirs):
        a1, a2 = _t(f1).requires_grad_(True), _t(f2).requires_grad_(True)
        blk = m.CorrBlock(a1[:, :2], a2[:, :2]).cat(m.CorrBlock(a1[:, 2:], a2[:, 2:]))
        out = blk(_t(co))
        (out * w).sum().backward()
        keep = torch.tensor([True, False, True, True]).cuda()
        sub = blk[keep](_t(co)[:, keep])
        res.append((out.detach(), a1.grad, a2.grad, sub.detach()))
    for x, y in zip(*res):
        assert torch.equal(_bits(x), _bits(y))


def test_reference_altcorrblock_rig2_lowmem(mods):
    """AltCorrBlock exactly as update_lowmem drives it: stereo buffer (rig = 2), f16 fmaps, chunk of edges incl. ii == jj."""
    ours, theirs = mods
    fmaps, ii, jj, co = D.altcorr_case(ht=48, wd=64)
    fm = _t(fmaps)
    num, rig, ch, ht, wd = fm.shape
    res = []
    for m in (ours, theirs):
        op = m.AltCorrBlock(fm.view(1, num * rig, ch, ht, wd))
        res.append(op(_t(co), _t(ii), _t(jj)))
    a, b = res
    assert a.shape == b.shape == (1, len(ii), 196, ht, wd) and a.dtype == b.dtype == torch.float32
    scale = float(b.abs().max())
    assert float((a - b).abs().max()) <= 1e-4 * scale, float((a - b).abs().max()) / scale
    assert float((a - b).norm() / b.norm()) <= 1e-5


def test_reference_altcorrblock_multi_sample_coords(mods):
    """the 6-D coords form [B,N,H,W,S,2] (corr.py:110-131) -> [B,N,196,H,W,S]"""
    ours, theirs = mods
    fmaps, ii, jj, co = D.altcorr_case(seed=44, ht=24, wd=32, S=3)
    fm = _t(fmaps)
    num, rig, ch, ht, wd = fm.shape
    a, b = [m.AltCorrBlock(fm.view(1, num * rig, ch, ht, wd))(_t(co), _t(ii), _t(jj)) for m in mods]
    assert a.shape == b.shape == (1, len(ii), 196, ht, wd, 3)
    assert float((a - b).abs().max()) <= 1e-4 * float(b.abs().max())


def test_reference_altcorr_backward_fp32(mods):
    """CorrLayer.backward through autograd: gradients reach the fp32 feature pyramid (training use of AltCorrBlock)."""
    ours, theirs = mods
    fmaps, ii, jj, co = D.altcorr_case(seed=45, ht=16, wd=24, dtype=np.float32)
    num, rig, ch, ht, wd = fmaps.shape
    w = _t(np.random.default_rng(2).normal(0, 1, (1, len(ii), 196, ht, wd)).astype(np.float32))
    res = []
    for m in (ours, theirs):
        fm = _t(fmaps).requires_grad_(True)
        out = m.AltCorrBlock(fm.view(1, num * rig, ch, ht, wd))(_t(co), _t(ii), _t(jj))
        (out * w).sum().backward()
        res.append((out.detach(), fm.grad))
    (oa, ga), (ob, gb) = res
    assert float((oa - ob).abs().max()) <= 1e-4 * float(ob.abs().max())
    assert float((ga - gb).norm() / gb.norm()) <= 1e-4
    assert float((ga - gb).abs().max()) <= 1e-4 * float(gb.abs().max())


# ------------------------------------------------------------------ the repo's own host mirror against the same reference run
def test_mirror_corrblock_f16(mods):
    """droid_b200.corr.CorrBlock (one fused pyramid build on tcgen05 + one fused 4-level lookup) against the reference
    classes on the reference extension.  Level 0 is the same f16 rounding of the same fp32 dot product as cuBLAS gives
    (<= 1 ulp where the accumulation order differs); levels 1-3 are computed as f1^T pool(f2) and rounded ONCE, while the
    reference rounds avg_pool2d of the rounded level below at every level — both are within f16 rounding of the exact
    value, they are not the same bits.  The achieved distance is asserted here in units of the f16 spacing at the value
    scale and recorded in DESIGN.md §3."""
    _, theirs = mods
    from droid_b200 import corr as mirror
    f1, f2, co = D.corrblock_case(ht=48, wd=64, n=3)
    a_blk = mirror.CorrBlock(_t(f1), _t(f2))
    b_blk = theirs.CorrBlock(_t(f1), _t(f2))
    worst = []
    for l in range(4):
        va, vb = a_blk.corr_pyramid[l].float(), b_blk.corr_pyramid[l].float()
        assert va.shape == vb.shape
        scale = float(vb.abs().max())
        ulp = 2.0 ** (np.floor(np.log2(scale)) - 10)                  # f16 spacing at the top of the value range
        worst.append(float((va - vb).abs().max()) / ulp)
    print("mirror CorrBlock f16 pyramid: max |diff| per level in f16 ulps at the value scale:", worst)
    assert worst[0] <= 1.0 and max(worst[1:]) <= 2.0, worst
    a, b = a_blk(_t(co)), b_blk(_t(co))
    assert a.shape == b.shape and a.dtype == b.dtype
    assert float((a.float() - b.float()).abs().max()) <= 4e-3 * float(b.float().abs().max())


def test_mirror_corrblock_tiled_levels_are_private():
    """The mirror's inference CorrBlock keeps levels 0 and 1 in the tiled slice layout: lookups, cat and indexing work on
    it directly; `corr_pyramid` hands out the reference's row-major volumes (converting the block once) and the lookup of
    the converted block gives the same bits."""
    from droid_b200 import corr as mirror
    f1, f2, co = D.corrblock_case(ht=48, wd=64, n=3)
    blk = mirror.CorrBlock(_t(f1), _t(f2))
    assert blk._tiled == 0b0011
    a = blk(_t(co))
    two = mirror.CorrBlock(_t(f1), _t(f2)).cat(mirror.CorrBlock(_t(f1), _t(f2)))
    assert two._tiled == 0b0011 and two._pyr[0].shape[0] == 6
    co2 = _t(np.concatenate([co, co], 1))
    aa = two(co2)
    assert torch.equal(_bits(aa[:, :3]), _bits(a)) and torch.equal(_bits(aa[:, 3:]), _bits(a))
    sel = two[torch.tensor([4, 0], device="cuda")]
    assert torch.equal(_bits(sel(co2[:, [4, 0]])), _bits(a[:, [1, 0]]))
    rows = blk.corr_pyramid                                          # converts: row-major from here on
    assert blk._tiled == 0 and [tuple(v.shape) for v in rows] == [(3, 48, 64, 48 >> l, 64 >> l) for l in range(4)]
    assert torch.equal(_bits(blk(_t(co))), _bits(a))
    fresh = mirror.CorrBlock(_t(f1), _t(f2))
    mixed = fresh.cat(blk)                                           # tiled + row-major: both end up row-major
    assert mixed._tiled == 0 and torch.equal(_bits(mixed(co2)), _bits(aa))


def test_mirror_corrblock_fp32_grads(mods):
    """the mirror's fused volume build must not drop the gradients to the feature maps."""
    _, theirs = mods
    from droid_b200 import corr as mirror
    f1, f2, co = D.corrblock_case(seed=48, dtype=np.float32, n=3, ht=16, wd=24)
    w = _t(np.random.default_rng(3).normal(0, 1, (1, 3, 196, 16, 24)).astype(np.float32))
    res = []
    for m in (mirror, theirs):
        a1, a2 = _t(f1).requires_grad_(True), _t(f2).requires_grad_(True)
        out = m.CorrBlock(a1, a2)(_t(co))
        (out * w).sum().backward()
        res.append((out.detach(), a1.grad, a2.grad))
    for x, y in zip(*res):
        assert x is not None and float((x - y).abs().max()) <= 1e-4 * float(y.abs().max())


def test_mirror_altcorrblock(mods):
    _, theirs = mods
    from droid_b200 import corr as mirror
    fmaps, ii, jj, co = D.altcorr_case(ht=48, wd=64)
    fm = _t(fmaps)
    num, rig, ch, ht, wd = fm.shape
    a = mirror.AltCorrBlock(fm.view(1, num * rig, ch, ht, wd))(_t(co), _t(ii), _t(jj))
    b = theirs.AltCorrBlock(fm.view(1, num * rig, ch, ht, wd))(_t(co), _t(ii), _t(jj))
    assert a.shape == b.shape and a.dtype == b.dtype
    assert float((a - b).abs().max()) <= 1e-4 * float(b.abs().max())


def test_empty_batches_return_empty_outputs():
    """no edges / no pairs: correctly shaped empty outputs and no launch (the reference launches a zero-block grid, which is a
    CUDA launch error) — through the nine-name operators and the mirror classes"""
    import droid_backends as B
    from droid_b200 import corr as mirror
    dev, H, W = "cuda", 24, 32
    vol = torch.zeros((0, H, W, H, W), device=dev, dtype=torch.float16)
    co = torch.zeros((0, 2, H, W), device=dev)
    assert B.corr_index_forward(vol, co, 3)[0].shape == (0, 7, 7, H, W)
    assert B.corr_index_backward(vol.float(), co, torch.zeros((0, 7, 7, H, W), device=dev), 3)[0].shape == vol.shape
    f, c5 = torch.zeros((0, H, W, 128), device=dev), torch.zeros((0, 1, H, W, 2), device=dev)
    assert B.altcorr_forward(f, f, c5, 3)[0].shape == (0, 1, 49, H, W)
    g1, g2, gc = B.altcorr_backward(f, f, c5, torch.zeros((0, 1, 49, H, W), device=dev), 3)
    assert g1.shape == f.shape and g2.shape == f.shape and gc.shape == c5.shape
    poses = torch.zeros((4, 7), device=dev); poses[:, 6] = 1
    e = torch.zeros(0, dtype=torch.int64, device=dev)
    assert B.frame_distance(poses, torch.ones((4, H, W), device=dev), torch.tensor([50., 50., 16., 12.], device=dev), e, e, 0.3).shape == (0,)
    fm = torch.randn((1, 6, 128, H, W), device=dev, dtype=torch.float16)
    assert mirror.AltCorrBlock(fm)(torch.zeros((1, 0, H, W, 2), device=dev), e, e).shape == (1, 0, 196, H, W)
    assert mirror.CorrBlock(fm[:, :0], fm[:, :0])(torch.zeros((1, 0, H, W, 2), device=dev)).shape[:2] == (1, 0)
    torch.cuda.synchronize()
