"""GPU parity of the tensor-core altcorr path (csrc/altcorr_tc.cu) against the CPU oracle (restatement of
/root/reference/src/altcorr_kernel.cu:27-149): the nine-name `altcorr_forward`, the fused pyramid entry point, and the
routes that leave the tensor-core kernel (cells beyond the 8x8 mask, values outside the fp16 range, NaN coordinates).
Tolerance (north_star): 1e-4 relative on fp32 correlation values = 1e-4 of the value scale."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import dropin_cases as D  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _check_status():
    import droid_backends
    old = droid_backends.CHECK_STATUS
    droid_backends.CHECK_STATUS = True          # every call reads the kernel's status words back: protocol errors fail loudly
    yield
    droid_backends.CHECK_STATUS = old


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _api_case(seed, B, S, H, W, H2, W2, exact_f16=False, scatter=False):
    rng = np.random.default_rng(seed)
    f1 = rng.normal(0, 1, (B, H, W, 128)).astype(np.float32)
    f2 = rng.normal(0, 1, (B, H2, W2, 128)).astype(np.float32)
    if exact_f16:
        f1, f2 = f1.astype(np.float16).astype(np.float32), f2.astype(np.float16).astype(np.float32)
    if scatter:
        co = np.stack([rng.uniform(-0.2 * W2, 1.2 * W2, (B, S, H, W)), rng.uniform(-0.2 * H2, 1.2 * H2, (B, S, H, W))], -1)
    else:
        co = np.stack([D.smooth_coords(rng, B, H, W) * np.array([W2 / W, H2 / H]) for _ in range(S)], 1)
    co = co.astype(np.float32)
    co[0, 0, 0, 0] = [3.0, 5.0]                      # exactly integer coordinates
    co[0, 0, 0, 1] = [-50.0, 1e4]                    # fully outside
    co[0, 0, 1, 0] = [W2 - 0.5, H2 - 0.5]            # window half outside at the far corner
    return f1, f2, co


@pytest.mark.parametrize("B,S,H,W,H2,W2,exact,scatter", [
    (2, 1, 48, 64, 48, 64, False, False),            # BASELINE map size, level 0, fp32 operands: 3 products
    (2, 1, 48, 64, 24, 32, True, False),             # level 1, fp16-exact operands: 1 product (lo == 0 detected at run time)
    (1, 2, 24, 32, 6, 8, False, False),              # level 3 of a small map, two coordinate samples
    (1, 1, 21, 29, 10, 14, False, False),            # ragged tiles (21 = 2*8 + 5 rows, 29 = 16 + 13 columns)
    (2, 1, 24, 32, 24, 32, False, True),             # uniformly scattered coordinates: every tile touches every cell
    (1, 1, 8, 16, 48, 64, True, True),               # one tile against a larger target map
])
def test_altcorr_forward_tc_vs_oracle(B, S, H, W, H2, W2, exact, scatter):
    import droid_backends
    from oracle import oracle
    f1, f2, co = _api_case(100 + H + W2, B, S, H, W, H2, W2, exact, scatter)
    ref = oracle.altcorr_forward(f1, f2, co, 3)
    out, = droid_backends.altcorr_forward(_t(f1), _t(f2), _t(co), 3)
    out = out.cpu().numpy()
    assert out.shape == ref.shape
    scale = np.sqrt(128.0)                           # magnitude of a dot product of N(0,1) features
    err = np.abs(out - ref).max()
    assert err <= 1e-4 * scale, err / scale
    assert np.linalg.norm(out - ref) / np.linalg.norm(ref) <= 2e-6


def test_altcorr_forward_cells_beyond_mask_go_to_simt_kernel():
    """A 72 x 160 target map with scattered coordinates: the windows of a tile span more than 8 x 8 cells (128 x 64
    positions) from their bounding corner -> those (tile, level) entries are handed to the SIMT kernel."""
    import droid_backends
    from droid_b200 import _capi
    from oracle import oracle
    f1, f2, co = _api_case(7, 1, 1, 16, 32, 72, 160, False, True)
    ref = oracle.altcorr_forward(f1, f2, co, 3)
    out, = droid_backends.altcorr_forward(_t(f1), _t(f2), _t(co), 3)
    assert np.abs(out.cpu().numpy() - ref).max() <= 1e-4 * np.sqrt(128.0)
    # the same call again, reading the status words: entries > 0
    lib = _capi.lib()
    B, S, H, W, H2, W2 = 1, 1, 16, 32, 72, 160
    wsb = lib.db200_altcorr_workspace_bytes(B, S, H, W, H2, W2, 128)
    ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
    o2 = torch.empty((B, S, 49, H, W), device="cuda")
    a, b, c = _t(f1), _t(f2), _t(co)
    import ctypes as C
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert lib.db200_altcorr_forward_ws(a.data_ptr(), b.data_ptr(), c.data_ptr(), o2.data_ptr(), B, S, H, W, H2, W2, 128, 3,
                                        ws.data_ptr(), wsb, st) == 0
    n = lib.db200_altcorr_status(C.c_void_p(ws.data_ptr() + 256), st)
    assert n > 0, n
    assert torch.equal(o2, out)


def test_altcorr_forward_out_of_fp16_range_and_nan_coords():
    import droid_backends
    from oracle import oracle
    f1, f2, co = _api_case(9, 1, 1, 16, 32, 16, 32)
    f1b = f1.copy(); f1b[0, 3, 4, 7] = 3.0e5         # outside the fp16 range: whole call on the SIMT kernel, fp32 operands
    ref = oracle.altcorr_forward(f1b, f2, co, 3)
    out, = droid_backends.altcorr_forward(_t(f1b), _t(f2), _t(co), 3)
    assert np.abs(out.cpu().numpy() - ref).max() <= 1e-5 * np.abs(ref).max()
    co2 = co.copy(); co2[0, 0, 5, 6] = [np.nan, 4.0]; co2[0, 0, 5, 7] = [np.inf, 4.0]
    out2, = droid_backends.altcorr_forward(_t(f1), _t(f2), _t(co2), 3)
    out2 = out2.cpu().numpy()
    ref2 = oracle.altcorr_forward(f1, f2, co, 3)
    assert np.isnan(out2[0, 0, :, 5, 6]).all() and np.isnan(out2[0, 0, :, 5, 7]).all()      # 0 * NaN weights, as in the reference
    m = np.ones(out2.shape, bool); m[0, 0, :, 5, 6:8] = False
    assert np.abs(out2[m] - ref2[m]).max() <= 1e-4 * np.sqrt(128.0)


@pytest.mark.parametrize("S", [None, 3])
@pytest.mark.parametrize("dtype", [np.float16, np.float32])
def test_altcorr_pyramid_forward_vs_oracle(S, dtype):
    """The fused entry point (all levels, frame gathers, casts, concat in one launch) against the oracle run the way
    AltCorrBlock.corr_fn runs the reference op: per level, on gathered fp32 copies, coords / 2**l."""
    import droid_backends
    from oracle import oracle
    from droid_b200 import corr as mirror
    fmaps, ii, jj, co = D.altcorr_case(seed=61, ht=48 if S is None else 24, wd=64 if S is None else 32, dtype=dtype, S=S)
    num, rig, ch, ht, wd = fmaps.shape
    blk = mirror.AltCorrBlock(_t(fmaps).view(1, num * rig, ch, ht, wd))
    assert blk._fused
    out = blk(_t(co), _t(ii), _t(jj)).cpu().numpy()
    E = len(ii)
    assert out.shape == ((1, E, 196, ht, wd) if S is None else (1, E, 196, ht, wd, S))
    pyr = [p[0].float().cpu().numpy() for p in blk.pyramid]
    co6 = co if S is not None else co[..., None, :]
    cs = np.ascontiguousarray(np.moveaxis(co6[0], 3, 1))                       # [E,S,H,W,2]
    for l in range(4):
        ref = oracle.altcorr_forward(pyr[0][ii], pyr[l][jj], (cs / 2 ** l).astype(np.float32), 3)    # [E,S,49,H,W]
        got = out[0, :, 49 * l:49 * (l + 1)]
        got = got[:, None] if S is None else np.moveaxis(got, -1, 1)
        assert np.abs(got - ref).max() <= 1e-4 * max(np.abs(ref).max(), 1e-3), l


@pytest.mark.parametrize("B,S,H,W,H2,W2,exact,scatter", [
    (2, 1, 48, 64, 48, 64, False, False),
    (2, 1, 24, 32, 12, 16, True, False),
    (1, 2, 21, 29, 10, 14, False, False),            # ragged tiles, two coordinate samples accumulate into the same gradients
    (1, 1, 16, 32, 24, 32, False, True),
])
def test_altcorr_backward_tc_vs_oracle(B, S, H, W, H2, W2, exact, scatter):
    import droid_backends
    from oracle import oracle
    f1, f2, co = _api_case(300 + H + W2, B, S, H, W, H2, W2, exact, scatter)
    g = np.random.default_rng(5).normal(0, 1, (B, S, 49, H, W)).astype(np.float32)
    r1, r2, _ = oracle.altcorr_backward(f1, f2, co, g, 3)
    g1, g2, gc = droid_backends.altcorr_backward(_t(f1), _t(f2), _t(co), _t(g), 3)
    for name, a, b in (("fmap1_grad", g1, r1), ("fmap2_grad", g2, r2)):
        a = a.cpu().numpy()
        assert np.abs(a - b).max() <= 1e-4 * np.abs(b).max(), (name, np.abs(a - b).max() / np.abs(b).max())
        assert np.linalg.norm(a - b) / np.linalg.norm(b) <= 1e-5, name
    assert float(gc.abs().max()) == 0.0 and gc.shape == (B, S, H, W, 2)


def test_altcorr_backward_handover_to_simt():
    """cells beyond the 8x8 mask (large scattered target map) and values outside the fp16 range: the SIMT kernel redoes the call"""
    import droid_backends
    from oracle import oracle
    f1, f2, co = _api_case(11, 1, 1, 16, 32, 72, 160, False, True)
    g = np.random.default_rng(6).normal(0, 1, (1, 1, 49, 16, 32)).astype(np.float32)
    for f1x in (f1, f1 * np.float32(3e4)):
        r1, r2, _ = oracle.altcorr_backward(f1x, f2, co, g, 3)
        g1, g2, _ = droid_backends.altcorr_backward(_t(f1x), _t(f2), _t(co), _t(g), 3)
        assert np.abs(g1.cpu().numpy() - r1).max() <= 1e-4 * np.abs(r1).max()
        assert np.abs(g2.cpu().numpy() - r2).max() <= 1e-4 * np.abs(r2).max()
