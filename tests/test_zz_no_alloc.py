"""The allocation-free LayerNorm-backward entry point, called through the C ABI.

``rt1_layernorm_bwd_noalloc`` exists for embedders that cannot give the library a scratch buffer — e.g. a control
process whose device memory was fully reserved at start-up, or a host whose memory is owned by another allocator, so
that neither they nor we may call ``cudaMalloc`` while the robot is running.  Its scratch is a fixed ``__device__``
array inside the library (max row width 1024), shared by every call on a device.

The test drives it with ``ctypes`` exactly as a C host would (raw pointers + stream) and checks it against the
regular entry point (caller-provided scratch) on two back-to-back calls with different data and shapes: the second
call must not be disturbed by what the first one left in the shared array.  This file sorts last on purpose, so a run
that stops at the first failure has already executed every other test.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from rt1b200.ops import _lib as loader
    return loader.lib()


def _p(t):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def _run_both(rows, D, seed):
    lib = _lib()
    dev = "cuda"
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(rows, D, generator=g) * 1.5 + 0.3).to(torch.bfloat16).to(dev)
    w = (torch.randn(D, generator=g) * 0.2 + 1.0).to(torch.bfloat16).to(dev)
    b = (torch.randn(D, generator=g) * 0.2).to(torch.bfloat16).to(dev)
    dy = torch.randn(rows, D, generator=g).to(torch.bfloat16).to(dev)
    y = torch.empty_like(x)
    mean = torch.empty(rows, dtype=torch.float32, device=dev)
    rstd = torch.empty_like(mean)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert lib.rt1_layernorm_fwd(_p(x), _p(w), _p(b), _p(y), _p(mean), _p(rstd), ctypes.c_int(rows), ctypes.c_int(D),
                                 ctypes.c_float(1e-5), stream) == 0
    # existing entry point: scratch handed in by the caller
    lib.rt1_layernorm_bwd_scratch_floats.restype = ctypes.c_longlong
    n_scratch = lib.rt1_layernorm_bwd_scratch_floats(ctypes.c_int(rows), ctypes.c_int(D))
    scratch = torch.empty(n_scratch, dtype=torch.float32, device=dev)
    dx_a, dw_a, db_a = torch.empty_like(x), torch.empty_like(w), torch.empty_like(b)
    assert lib.rt1_layernorm_bwd(_p(x), _p(w), _p(dy), _p(mean), _p(rstd), _p(dx_a), _p(dw_a), _p(db_a), _p(scratch),
                                 ctypes.c_int(rows), ctypes.c_int(D), stream) == 0
    # new entry point: nothing to allocate, nothing to pass
    dx_b, dw_b, db_b = torch.empty_like(x), torch.empty_like(w), torch.empty_like(b)
    rc = lib.rt1_layernorm_bwd_noalloc(_p(x), _p(w), _p(dy), _p(mean), _p(rstd), _p(dx_b), _p(dw_b), _p(db_b),
                                       ctypes.c_int(rows), ctypes.c_int(D), stream)
    assert rc == 0, lib.rt1_last_error()
    torch.cuda.synchronize()
    return (dx_a, dw_a, db_a), (dx_b, dw_b, db_b)


def test_noalloc_layernorm_bwd_matches_regular_entry_point_twice_in_a_row():
    # first call on this device also resolves the scratch symbol's device address (done outside any stream capture)
    for rows, D, seed in ((528, 512, 0), (1200, 1024, 1)):        # second call: more CTAs, wider rows, other data
        ref, got = _run_both(rows, D, seed)
        for name, a, b in zip(("dx", "dgamma", "dbeta"), ref, got):
            torch.testing.assert_close(b.float(), a.float(), rtol=2e-2, atol=2e-2, msg=lambda m: f"{name} rows={rows} D={D}: {m}")
    # and a smaller problem afterwards, which leaves most of the stale array untouched
    ref, got = _run_both(40, 256, 2)
    for a, b in zip(ref, got):
        torch.testing.assert_close(b.float(), a.float(), rtol=2e-2, atol=2e-2)


def test_noalloc_rejects_rows_wider_than_the_built_in_scratch():
    lib = _lib()
    lib.rt1_last_error.restype = ctypes.c_char_p
    lib.rt1_layernorm_bwd_noalloc_max_d.restype = ctypes.c_longlong
    max_d = lib.rt1_layernorm_bwd_noalloc_max_d()
    assert max_d == 1024
    t = torch.zeros(8, 2048, dtype=torch.bfloat16, device="cuda")
    v = torch.zeros(2048, dtype=torch.bfloat16, device="cuda")
    s = torch.zeros(8, dtype=torch.float32, device="cuda")
    rc = lib.rt1_layernorm_bwd_noalloc(_p(t), _p(v), _p(t), _p(s), _p(s), _p(t), _p(v), _p(v), ctypes.c_int(8),
                                       ctypes.c_int(2048), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    assert b"exceeds the built-in scratch" in lib.rt1_last_error()
