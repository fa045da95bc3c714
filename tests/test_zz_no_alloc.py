"""fa_b200_bwd_noalloc (``-m gpu``): the backward whose workspace is the library's static device array.

Collected last on purpose (the file name sorts after every other test file):
a failure here cannot hide another test.  The entry point is called through the
C ABI (ctypes, raw pointers) and compared with the existing entry point, fa_b200_bwd with a
caller-provided workspace, on the same inputs -- several calls in a row, of different sizes, so that
every call after the first finds the previous call's contents in the shared scratch.

Tolerance: the two entry points run the same kernels, so dK / dV must be bit-identical; dQ is summed
by fp32 reduce-adds in arrival order (~1e-7 relative from run to run, which can flip the last bit of a
16-bit result), so it gets the relative Frobenius bound the other gradient tests use (1e-2) and, in
practice, lands several orders below it (asserted at 1e-3).
"""
import ctypes

import pytest
import torch

from conftest import SEED

pytestmark = pytest.mark.gpu


def _native():
    from flash_attention_ipu_b200._impl import native
    assert native.library_path().endswith("libflash_attention.so")
    return native


def _rel_fro(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30)).item()


def _case(native, dtype, B, H, Hkv, Lq, Lk, D, causal, q0, seed):
    g = torch.Generator().manual_seed(seed)
    q, dout = (torch.randn(B, H, Lq, D, generator=g).to(dtype).cuda() for _ in range(2))
    k, v = (torch.randn(B, Hkv, Lk, D, generator=g).to(dtype).cuda() for _ in range(2))
    prob = native.make_problem(q, k, causal, D ** -0.5, 0, 0, q0, 0)
    o = torch.empty_like(q)
    lse = torch.empty(B, H, Lq, device="cuda", dtype=torch.float32)
    native.fwd(prob, q, k, v, o, lse)
    return prob, (q, k, v, o, dout, lse)


def _with_workspace(native, prob, ops):
    q, k, v, o, dout, lse = ops
    dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
    ws = torch.empty(max(native.bwd_workspace_bytes(prob), 256), device="cuda", dtype=torch.uint8)
    native.bwd(prob, q, k, v, o, dout, lse, dq, dk, dv, ws)
    torch.cuda.synchronize()
    return dq, dk, dv


def _no_alloc(native, prob, ops, outs=None):
    q, k, v, o, dout, lse = ops
    # outputs belong to the caller and are allocated BEFORE the call; NaN-filled so that anything the
    # kernels fail to write shows up
    dq, dk, dv = outs or tuple(torch.full_like(t, float("nan")) for t in (q, k, v))
    native.bwd_noalloc(prob, q, k, v, o, dout, lse, dq, dk, dv)
    torch.cuda.synchronize()
    return dq, dk, dv


def _check(got, want, what):
    dq, dk, dv = got
    rq, rk, rv = want
    assert torch.isfinite(dq.float()).all() and torch.isfinite(dk.float()).all() and torch.isfinite(dv.float()).all(), what
    assert torch.equal(dk, rk), f"{what}: dK differs from fa_b200_bwd"
    assert torch.equal(dv, rv), f"{what}: dV differs from fa_b200_bwd"
    err = _rel_fro(dq, rq)
    assert err < 1e-2 and err < 1e-3, f"{what}: dQ relative error {err}"


def test_bwd_noalloc_matches_bwd_on_calls_in_a_row():
    native = _native()
    # (dtype, B, H, Hkv, Lq, Lk, D, causal, q_pos0): big tensor-family problem first, then smaller ones
    # (other head dim / dtype / grouped-query / ragged lengths / shifted diagonal), then the SIMT family
    cases = [
        (torch.bfloat16, 1, 4, 4, 1024, 1024, 128, True, 0),
        (torch.float16, 2, 4, 2, 200, 328, 64, True, 128),
        (torch.bfloat16, 1, 2, 2, 384, 384, 128, False, 0),
        (torch.float32, 1, 2, 2, 96, 96, 40, True, 0),
    ]
    built = [_case(native, *c, seed=SEED + i) for i, c in enumerate(cases)]
    assert native.kernel_family(built[0][0]) == 1 and native.kernel_family(built[3][0]) == 0
    want = [_with_workspace(native, prob, ops) for prob, ops in built]
    for rnd in range(2):  # every problem twice: the second round starts from the last problem's leftovers
        for i, (prob, ops) in enumerate(built):
            assert native.bwd_workspace_bytes(prob) <= native.NOALLOC_SCRATCH_BYTES
            before = native.launch_count()
            got = _no_alloc(native, prob, ops)
            assert native.launch_count() > before  # the library's kernels ran
            _check(got, want[i], f"round {rnd}, case {i}")


def test_bwd_noalloc_allocates_nothing_and_refuses_what_does_not_fit():
    native = _native()
    prob, ops = _case(native, torch.bfloat16, 1, 8, 8, 2048, 2048, 128, True, 0, seed=SEED + 7)
    need = native.bwd_workspace_bytes(prob)
    assert 8 << 20 < need <= native.NOALLOC_SCRATCH_BYTES
    want = _with_workspace(native, prob, ops)
    outs = tuple(torch.full_like(t, float("nan")) for t in (ops[0], ops[1], ops[2]))
    _check(_no_alloc(native, prob, ops, outs), want, "first call")  # modules, kernels and attributes are now set up
    for t in outs:
        t.fill_(float("nan"))
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    reserved0 = torch.cuda.memory_reserved()
    got = _no_alloc(native, prob, ops, outs)
    free1, _ = torch.cuda.mem_get_info()
    # a cudaMalloc of this problem's workspace (> 8 MiB) would lower the device's free memory
    assert free1 >= free0 - (1 << 20), (free0, free1)
    assert torch.cuda.memory_reserved() == reserved0
    _check(got, want, "second call")

    # beyond the static scratch: refused before anything is read or launched, with a clear message
    q = ops[0]
    big = native.Problem(4, 16, 8192, 8192, 128, native.FA_B200_BF16, 1, 0.0883883, 0, 0, 0, 0, 0, None)
    assert native.bwd_workspace_bytes(big) > native.NOALLOC_SCRATCH_BYTES
    t = native.Tensor(q.data_ptr(), 16 * 8192 * 128, 8192 * 128, 128)
    before = native.launch_count()
    rc = native._lib.fa_b200_bwd_noalloc(ctypes.byref(big), *([ctypes.byref(t)] * 5), ops[5].data_ptr(),
                                         *([ctypes.byref(t)] * 3), None)
    assert rc == -4  # FA_B200_ERR_WORKSPACE
    msg = native._lib.fa_b200_last_error().decode()
    assert "static scratch" in msg and "FA_B200_NOALLOC_SCRATCH_BYTES" in msg
    assert native.launch_count() == before
    with pytest.raises(RuntimeError, match="static scratch"):
        native._check(rc, "fa_b200_bwd_noalloc")
