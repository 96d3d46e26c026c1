"""rg_norm_stats_static -- the allocation-free variant of rg_norm_stats (no counterpart in the reference; added
for callers that cannot call cudaMalloc).  Its partial-row scratch is a library-owned
`static __device__` array whose address the library resolves with cudaGetSymbolAddress once per device.

Called through the C ABI (ctypes) and checked against the existing entry point, which takes a caller workspace, on two calls
in a row (the second sees the first call's leftovers in the shared array); shapes beyond the built-in maximum are refused
with a message.  This file is collected LAST and passes alone in a fresh process as well as
after the other tests."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
MAX_FLOATS = 1 << 20          # RG_NORM_STATS_STATIC_MAX_FLOATS (include/raygun_b200.h)


def _stats(fn_name, x, kind, ws=None):
    from raygun_b200 import _C
    lib = _C.lib()
    groups = x.shape[0] * x.shape[1] if kind == _C.NORM_INSTANCE else x.shape[1]
    out = torch.full((2, groups), float("nan"), dtype=torch.float32, device=x.device)
    d = _C.t5(x)
    stream = C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)
    args = [C.byref(d), C.c_void_p(x.data_ptr()), C.c_int(kind), C.c_void_p(out[0].data_ptr()), C.c_void_p(out[1].data_ptr())]
    if ws is not None:
        args.append(C.c_void_p(ws.data_ptr()))
    rc = getattr(lib, fn_name)(*args, stream)
    return rc, out


def test_static_scratch_entry_point_matches_the_workspace_entry_point_on_consecutive_calls():
    from raygun_b200 import _C
    from raygun_b200 import kernels as K
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)                       # the entry point resolves the scratch of the CURRENT device
    _C.require_device(torch.empty(1, device=dev))
    gen = torch.Generator().manual_seed(5)
    cases = [
        # dense NDHWC bf16 (what the convolutions write), instance statistics
        (torch.randn(2, 64, 12, 10, 14, generator=gen).to(dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last_3d),
         _C.NORM_INSTANCE),
        # fp32 NCDHW, batch statistics, different shape: the second call reuses the array the first one left dirty
        (torch.randn(3, 40, 7, 9, 11, generator=gen).to(dev) * 3 + 1, _C.NORM_BATCH),
        # a strided NDHWC view (pad-arena interior), instance statistics
        (torch.randn(1, 128, 10, 12, 12, generator=gen).to(dev).to(torch.bfloat16).contiguous(
            memory_format=torch.channels_last_3d)[:, :, 1:-1, 1:-1, 1:-1], _C.NORM_INSTANCE),
    ]
    for rounds in range(2):                          # the whole sequence twice: every shape follows every other
        for x, kind in cases:
            ws = K.reduce_ws(x.shape[0], x.shape[1], dev)
            rc_ref, ref = _stats("rg_norm_stats", x, kind, ws)
            rc_new, got = _stats("rg_norm_stats_static", x, kind)
            torch.cuda.synchronize()
            assert rc_ref == 0 and rc_new == 0, _C.lib().rg_last_error()
            assert torch.isfinite(got).all()
            # same kernels, same partial-row order: identical bits (the GPU tests' tolerance for this op is 1e-4 relative)
            assert torch.equal(got, ref), float((got - ref).abs().max())
            xs = x.float()
            dims = (2, 3, 4) if kind == _C.NORM_INSTANCE else (0, 2, 3, 4)
            want = xs.sum(dim=dims).flatten()
            assert float((got[0] - want).abs().max()) <= 1e-4 * max(1.0, float(want.abs().max())) * 10


def test_static_scratch_entry_point_refuses_shapes_beyond_its_maximum():
    from raygun_b200 import _C
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    lib = _C.lib()
    n, c_ok, c_big = 2, 439, 440
    assert int(lib.rg_reduce_ws_floats(C.c_longlong(n), C.c_longlong(c_ok))) <= MAX_FLOATS
    assert int(lib.rg_reduce_ws_floats(C.c_longlong(n), C.c_longlong(c_big))) > MAX_FLOATS
    x = torch.randn(n, c_big, 2, 2, 2, device=dev)
    rc, _ = _stats("rg_norm_stats_static", x, _C.NORM_INSTANCE)
    msg = lib.rg_last_error().decode()
    assert rc == -1 and "scratch" in msg and str(MAX_FLOATS) in msg, (rc, msg)
    x = torch.randn(n, c_ok, 2, 2, 2, device=dev)
    rc, got = _stats("rg_norm_stats_static", x, _C.NORM_INSTANCE)
    torch.cuda.synchronize()
    assert rc == 0, lib.rg_last_error()
    want = x.sum(dim=(2, 3, 4)).flatten()
    assert float((got[0] - want).abs().max()) <= 1e-4 * max(1.0, float(want.abs().max()))
