"""Allocation-free LayerNorm backward (``dtb_layernorm_bwd_noalloc``), called through the C ABI.

Who this entry point is for: hosts that drive ``libdtb200.so`` directly and have no way to obtain a
scratch buffer — a C/Fortran program without a device allocator, a plugin living inside a framework
that has reserved the whole GPU and does not let guests call ``cudaMalloc``, or a loop that cannot
afford ``cudaMalloc``'s device-wide synchronisation.  Its scratch is a fixed array inside the library,
so the only thing to verify is that it computes exactly what ``dtb_layernorm_bwd`` computes with a
caller-provided buffer — including on a second call, when the built-in array still holds the first
call's partial sums.

The file name sorts last on purpose: under ``pytest -x`` every other test has run before this one.
It must also pass on its own in a fresh process (the first call resolves the array's device address).
"""
import ctypes as C

import pytest
import torch

from distributed_training_b200.ops import native

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _ptr(t):
    return C.c_void_p(t.data_ptr() if t is not None else 0)


def _inputs(rows, cols, seed):
    g = torch.Generator().manual_seed(seed)
    mk = lambda *s: torch.randn(*s, generator=g).to(torch.bfloat16).to(DEV)  # noqa: E731
    x, dy, dres = mk(rows, cols), mk(rows, cols), mk(rows, cols)
    gamma, beta = mk(cols) * 0.1 + 1, mk(cols) * 0.1
    y = torch.empty_like(x)
    mean, rstd = torch.empty(rows, device=DEV), torch.empty(rows, device=DEV)
    native.layernorm_fwd(x, None, gamma, beta, y, None, mean, rstd)
    return x, dy, dres, gamma, mean, rstd


def _run_both(rows, cols, seed):
    lib = native.lib()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    x, dy, dres, gamma, mean, rstd = _inputs(rows, cols, seed)
    # existing entry point, caller-provided scratch
    dx_a = torch.empty_like(x)
    dg_a, db_a = torch.empty(cols, device=DEV), torch.empty(cols, device=DEV)
    scratch = torch.empty(native.layernorm_bwd_scratch_floats(rows, cols), device=DEV)
    rc = lib.dtb_layernorm_bwd(_ptr(dy), _ptr(x), _ptr(gamma), _ptr(mean), _ptr(rstd), _ptr(dres), _ptr(dx_a),
                               _ptr(dg_a), _ptr(db_a), _ptr(scratch), rows, cols, 0, stream)
    assert rc == 0, lib.dtb_last_error().decode()
    # allocation-free entry point: inputs and outputs only
    dx_b = torch.empty_like(x)
    dg_b, db_b = torch.empty(cols, device=DEV), torch.empty(cols, device=DEV)
    rc = lib.dtb_layernorm_bwd_noalloc(_ptr(dy), _ptr(x), _ptr(gamma), _ptr(mean), _ptr(rstd), _ptr(dres), _ptr(dx_b),
                                       _ptr(dg_b), _ptr(db_b), rows, cols, 0, stream)
    assert rc == 0, lib.dtb_last_error().decode()
    torch.cuda.synchronize()
    return (dx_a, dg_a, db_a), (dx_b, dg_b, db_b)


def test_noalloc_matches_scratch_entry_point_on_two_consecutive_calls():
    # second call: different shape and data, so leftovers of the first call in the built-in array would show
    for rows, cols, seed in [(2048, 768, 1), (300, 1600, 2)]:
        (dx_a, dg_a, db_a), (dx_b, dg_b, db_b) = _run_both(rows, cols, seed)
        assert torch.allclose(dx_b.float(), dx_a.float(), atol=6e-2, rtol=3e-2)
        assert torch.allclose(dg_b, dg_a, atol=1e-4, rtol=1e-5)
        assert torch.allclose(db_b, db_a, atol=1e-4, rtol=1e-5)
        assert torch.isfinite(dg_b).all() and torch.isfinite(db_b).all()


def test_noalloc_rejects_rows_wider_than_builtin_scratch():
    lib = native.lib()
    cols = 4096                                      # > DTB_LN_NOALLOC_MAX_COLS (2048)
    z = torch.zeros(8, cols, dtype=torch.bfloat16, device=DEV)
    f = torch.zeros(cols, device=DEV)
    rc = lib.dtb_layernorm_bwd_noalloc(_ptr(z), _ptr(z), _ptr(z[0]), _ptr(f), _ptr(f), None, _ptr(z), _ptr(f), _ptr(f),
                                       8, cols, 0, C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    msg = lib.dtb_last_error().decode()
    assert "2048" in msg and "scratch" in msg
