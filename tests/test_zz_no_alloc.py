"""db200_altcorr_forward_static — the no-allocation variant of altcorr_forward (scratch in a static device array inside
libdroid_b200.so; no counterpart in the reference).  Called through the
C-ABI and compared with the existing entry point db200_altcorr_forward_ws on two calls in a row: the second call has
other sizes and fp16-exact operands, so stale scratch contents of the first (lo parts, the "lo != 0" flag, status words)
would show.  Collected last on purpose.  Tolerance: 1e-4 of the value scale, as in
tests/test_gpu_altcorr.py (the two entry points run the same kernels: in practice the bits agree)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(seed, B, S, H, W, H2, W2, exact_f16):
    rng = np.random.default_rng(seed)
    f1 = rng.normal(0, 1, (B, H, W, 128)).astype(np.float32)
    f2 = rng.normal(0, 1, (B, H2, W2, 128)).astype(np.float32)
    if exact_f16:
        f1, f2 = f1.astype(np.float16).astype(np.float32), f2.astype(np.float16).astype(np.float32)
    yy, xx = np.meshgrid(np.arange(H, dtype=np.float64), np.arange(W, dtype=np.float64), indexing="ij")
    co = np.stack([xx * W2 / W, yy * H2 / H], -1)[None, None] + rng.normal(0, 2.5, (B, S, H, W, 2))
    co[0, 0, 0, 0] = [-40.0, 3.0]                     # a window fully outside the image
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a.astype(np.float32))).cuda()
    return t(f1), t(f2), t(co)


def _run_both(lib, f1, f2, co):
    B, S, H, W, _ = co.shape
    _, H2, W2, _ = f2.shape
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda x: C.c_void_p(x.data_ptr())
    # existing entry point: caller-provided workspace
    wsb = lib.db200_altcorr_workspace_bytes(B, S, H, W, H2, W2, 128)
    ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
    ref = torch.full((B, S, 49, H, W), 777.0, device="cuda")
    assert lib.db200_altcorr_forward_ws(p(f1), p(f2), p(co), p(ref), B, S, H, W, H2, W2, 128, 3, p(ws), wsb, st) == 0
    # new entry point: inputs and outputs only
    out = torch.full((B, S, 49, H, W), -777.0, device="cuda")
    assert wsb <= lib.db200_altcorr_static_max_bytes()
    assert lib.db200_altcorr_forward_static(p(f1), p(f2), p(co), p(out), B, S, H, W, H2, W2, 128, 3, st) == 0
    torch.cuda.synchronize()
    return out, ref


def test_altcorr_forward_static_matches_workspace_entry_point_twice_in_a_row():
    from droid_b200 import _capi
    lib = _capi.lib()
    assert lib.db200_altcorr_static_max_bytes() == 8 << 20
    scale = float(np.sqrt(128.0))
    # call 1: fp32 operands (hi + lo products, flag word set), two edges at the BASELINE map size
    out, ref = _run_both(lib, *_case(1, 2, 1, 48, 64, 48, 64, False))
    assert not bool((out == -777.0).any())
    assert float((out - ref).abs().max()) <= 1e-4 * scale
    # call 2, right after: other sizes, fp16-exact operands (lo == 0: the flag of call 1 must not survive), 2 samples
    out2, ref2 = _run_both(lib, *_case(2, 1, 2, 24, 32, 12, 16, True))
    assert not bool((out2 == -777.0).any())
    assert float((out2 - ref2).abs().max()) <= 1e-4 * scale
    # and call 1's problem once more: same result as before
    out3, _ = _run_both(lib, *_case(1, 2, 1, 48, 64, 48, 64, False))
    assert torch.equal(out3, out)


def test_altcorr_forward_static_rejects_what_does_not_fit():
    from droid_b200 import _capi
    lib = _capi.lib()
    B, S, H, W = 8, 1, 48, 64                          # needs ~25 MB of scratch: beyond the 8 MiB static array
    assert lib.db200_altcorr_workspace_bytes(B, S, H, W, H, W, 128) > lib.db200_altcorr_static_max_bytes()
    f1 = torch.zeros((B, H, W, 128), device="cuda"); co = torch.zeros((B, S, H, W, 2), device="cuda")
    out = torch.full((B, S, 49, H, W), 5.0, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda x: C.c_void_p(x.data_ptr())
    rc = lib.db200_altcorr_forward_static(p(f1), p(f1), p(co), p(out), B, S, H, W, H, W, 128, 3, st)
    assert rc == -6 and b"static scratch" in lib.db200_status_string(rc)
    torch.cuda.synchronize()
    assert bool((out == 5.0).all())                    # nothing was launched
    with pytest.raises(RuntimeError):
        _capi.check(rc, "altcorr_forward_static")
    # unsupported shape (C != 128): a clear argument error, not a silent other path
    assert lib.db200_altcorr_forward_static(p(f1), p(f1), p(co), p(out), 1, 1, 8, 16, 8, 16, 64, 3, st) == -1
