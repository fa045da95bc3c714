"""GPU parity tests (``-m gpu``): the CUDA path, called through the public API / C ABI, against
the CPU oracle on the same seeded inputs and against the golden fixtures produced by the
reference's own code.  Mirrors tests/python/test_flash_attention_qkv_packed.py:56-93 and
tests/python/test_scaled_dot_product_attention.py:59-153 of the reference with device="cuda".

Tolerances (stated here):
  * fp32 inputs: atol 1e-3 / rtol 1e-5 -- the reference's own fp32 row (…_packed.py:78).
  * 16-bit forward: max-abs <= 1e-2 against the fp32 oracle on the same rounded inputs.
  * 16-bit gradients: max-abs within 2x (max-abs over ~1e6 elements is a noisy statistic) of the
    error of the reference math evaluated in the same dtype on
    the same inputs (north_star's second clause; bf16 cannot reach 1e-2 max-abs on N(0,1)
    data even in torch's own kernels), never above the reference's accepted 1.0
    (tests/cpp/test_flash_attention.cpp:292-295), plus a relative Frobenius bound of 1e-2.
"""
import numpy as np
import pytest
import torch

from conftest import GOLDEN, SEED, packed_inputs, sdpa_inputs
from oracle import attention_oracle as ao

pytestmark = pytest.mark.gpu
DT = {"bf16": torch.bfloat16, "fp16": torch.float16, "fp32": torch.float32}


def _native_loaded():
    from flash_attention_ipu_b200._impl import native
    assert native.library_path().endswith("libflash_attention.so")
    return native


def run_packed(qkv_cpu, grad_cpu, dtype, ncq=1, nckv=1, **kw):
    import flash_attention_ipu as fa
    x = qkv_cpu.to("cuda", dtype).requires_grad_(True)
    out = fa.flash_attention_qkv_packed(x, ncq, nckv, **kw)
    out.backward(grad_cpu.to("cuda", dtype))
    torch.cuda.synchronize()
    return out.detach().float().cpu(), x.grad.float().cpu()


def rel_fro(a, b):
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def same_dtype_reference_error(qkv, grad, dtype, causal=True):
    """max-abs error of the reference's math run in `dtype` on the CPU (per q/k/v gradient)."""
    x = qkv.to(dtype).requires_grad_(True)
    (ao.packed_attention_reference(x, causal) * grad.to(dtype)).sum().backward()
    x32 = qkv.to(dtype).float().requires_grad_(True)
    (ao.packed_attention_reference(x32, causal) * grad.to(dtype).float()).sum().backward()
    return (x.grad.float() - x32.grad).abs().amax(dim=(1, 2, 3))


@pytest.mark.parametrize("name,G,L,D", [("packed_g2_l256_d128", 2, 256, 128), ("packed_g3_l128_d64", 3, 128, 64),
                                        ("packed_g4_l6_d2", 4, 6, 2)])
@pytest.mark.parametrize("dtype", ["fp32", "bf16", "fp16"])
def test_golden_fixtures(name, G, L, D, dtype):
    """CUDA path vs outputs of the reference's own code (tests/golden, oracle/gen_golden.py)."""
    _native_loaded()
    gold = np.load(GOLDEN / f"{name}.npz")
    qkv, grad = packed_inputs(G, L, D)
    out, dqkv = run_packed(qkv, grad, DT[dtype])
    gout, gd = torch.from_numpy(gold["out"]), torch.from_numpy(gold["dqkv"])
    if dtype == "fp32":
        torch.testing.assert_close(out, gout, atol=1e-3, rtol=1e-5)
        torch.testing.assert_close(dqkv, gd, atol=1e-3, rtol=1e-5)
    else:
        # golden is for unrounded inputs: allow input-rounding noise on top (reference row: 1e-2/1e-2)
        torch.testing.assert_close(out, gout, atol=3e-2, rtol=1e-2)
        assert rel_fro(dqkv, gd) < 2e-2


@pytest.mark.parametrize("L", [256, 1024, 4096])
@pytest.mark.parametrize("dtype", ["bf16", "fp16", "fp32"])
def test_packed_vs_oracle(L, dtype):
    """reference matrix: N=4, D=128, chunks 16/16 (test_flash_attention_qkv_packed.py:56-76)."""
    G, D = 4, 128
    qkv, grad = packed_inputs(G, L, D)
    qkv, grad = qkv.to(DT[dtype]).float(), grad.to(DT[dtype]).float()  # oracle sees the rounded inputs
    out, dqkv = run_packed(qkv, grad, DT[dtype], 16, 16)
    ref = ao.packed_attention_reference(qkv)
    dref = ao.vanilla_attention_grad(grad, qkv)
    if dtype == "fp32":
        torch.testing.assert_close(out, ref, atol=1e-3, rtol=1e-5)
        torch.testing.assert_close(dqkv, dref, atol=1e-3, rtol=1e-5)
        return
    assert (out - ref).abs().max().item() <= 1e-2
    err = (dqkv - dref).abs().amax(dim=(1, 2, 3))
    bound = torch.maximum(2.0 * same_dtype_reference_error(qkv, grad, DT[dtype]), torch.tensor(1e-2))
    assert (err <= bound).all() and (err <= 1.0).all(), (err, bound)
    for t in range(3):
        assert rel_fro(dqkv[t], dref[t]) < 1e-2


@pytest.mark.parametrize("dtype", ["bf16", "fp16"])
def test_packed_long_sequence(dtype):
    """L=16384 row of the reference matrix (G reduced to 1 so the CPU oracle stays in seconds);
    oracle = the blocked restatement of the chunked algorithm."""
    G, L, D = 1, 16384, 128
    qkv, grad = packed_inputs(G, L, D)
    qkv, grad = qkv.to(DT[dtype]).float(), grad.to(DT[dtype]).float()
    out, dqkv = run_packed(qkv, grad, DT[dtype], 16, 16)
    ref, _ = ao.chunked_attention_forward(qkv, 16, 16)
    dref = ao.chunked_attention_backward(grad, qkv, 16, 16)
    assert (out - ref).abs().max().item() <= 1e-2
    for t in range(3):
        assert rel_fro(dqkv[t], dref[t]) < 1e-2
        assert (dqkv[t] - dref[t]).abs().max().item() <= 1.0


@pytest.mark.parametrize("dtype", ["bf16", "fp32"])
def test_chunk_hints_do_not_change_the_result(dtype):
    """tests/cpp/test_flash_attention.cpp:245-259: every chunking gives the same answer."""
    qkv, grad = packed_inputs(2, 512, 64)
    base = run_packed(qkv, grad, DT[dtype], 1, 1)
    for ncq, nckv in [(2, 2), (4, 2), (2, 4), (8, 8), (512, 256)]:
        got = run_packed(qkv, grad, DT[dtype], ncq, nckv)
        assert torch.equal(got[0], base[0])
        # dQ accumulates by fp32 reduce-adds in arrival order: a sum may round to the neighbouring 16-bit value
        # (one such flip among 65 536 elements is already 1.5e-5 relative Frobenius)
        assert rel_fro(got[1], base[1]) < 1e-3


@pytest.mark.parametrize("G,L,D", [(2, 200, 128), (1, 8, 64), (3, 331, 64), (1, 129, 128), (2, 1, 64)])
@pytest.mark.parametrize("causal", [True, False])
def test_ragged_lengths_and_noncausal(G, L, D, causal):
    """edge cases: lengths that are not tile multiples, single row, non-causal extension (unpinned)."""
    qkv, grad = packed_inputs(G, L, D)
    qkv, grad = qkv.bfloat16().float(), grad.bfloat16().float()
    out, dqkv = run_packed(qkv, grad, torch.bfloat16, causal=causal)
    ref = ao.packed_attention_reference(qkv, causal)
    dref = ao.vanilla_attention_grad(grad, qkv, causal)
    assert (out - ref).abs().max().item() <= 1e-2
    assert rel_fro(dqkv, dref) < 1e-2


def test_causality_property_full_size():
    """size-independent property at a BASELINE config-3 slice (T=8192, d=128): the first n rows
    of the output depend only on the first n rows of q, k, v."""
    import flash_attention_ipu as fa
    torch.manual_seed(SEED)
    qkv = torch.randn(3, 2, 8192, 128, device="cuda", dtype=torch.bfloat16)
    qkv[0] *= 128 ** -0.5
    full = fa.flash_attention_qkv_packed(qkv, 8, 8)
    for n in (128, 1000, 4096):
        part = fa.flash_attention_qkv_packed(qkv[:, :, :n].contiguous(), 1, 1)
        assert torch.equal(part, full[:, :n])
    # row 0 attends only to key 0 -> out[0] == v[0]
    assert torch.equal(full[:, 0], qkv[2][:, 0])
    # linearity in v: attn(q, k, 2v) == 2 attn(q, k, v) exactly (power-of-two scaling)
    qkv2 = qkv.clone(); qkv2[2] *= 2
    assert torch.equal(fa.flash_attention_qkv_packed(qkv2, 8, 8), 2 * full)


@pytest.mark.parametrize("batch", [(2,), (2, 3), (2, 3, 5)])
@pytest.mark.parametrize("L", [256, 1024])
@pytest.mark.parametrize("dtype", ["fp32", "fp16", "bf16"])
def test_sdpa_override(batch, L, dtype):
    """tests/python/test_scaled_dot_product_attention.py:59-113: patched SDPA on the device vs
    torch CPU SDPA, out and grad_q/k/v."""
    import torch.nn.functional as F
    import flash_attention_ipu_b200._impl.scaled_dot_product_attention as mod
    native = _native_loaded()
    qkv, grad = sdpa_inputs(batch, L, 64)
    qkv, grad = qkv.to(DT[dtype]).float(), grad.to(DT[dtype]).float()
    xr = qkv.clone().requires_grad_(True)
    ref = ao.sdpa_reference(xr[0], xr[1], xr[2])
    (ref * grad).sum().backward()
    mod.install()
    try:
        before = native.launch_count()
        x = qkv.to("cuda", DT[dtype]).requires_grad_(True)
        out = F.scaled_dot_product_attention(x[0], x[1], x[2], is_causal=True)
        out.backward(grad.to("cuda", DT[dtype]))
        torch.cuda.synchronize()
        assert native.launch_count() > before  # the native kernels ran, not torch's
    finally:
        mod.uninstall()
    out, g = out.detach().float().cpu(), x.grad.float().cpu()
    assert out.shape == ref.shape
    if dtype == "fp32":
        torch.testing.assert_close(out, ref.detach(), atol=1e-3, rtol=1e-5)
        torch.testing.assert_close(g, xr.grad, atol=1e-3, rtol=1e-5)
    else:
        assert (out - ref.detach()).abs().max().item() <= 1e-2
        assert rel_fro(g, xr.grad) < 1e-2


def test_sdpa_strided_nanogpt_layout():
    """q, k, v as nanoGPT makes them (demo/nanoGPT/model.py:56-59): (B,T,3C) split, viewed
    (B,T,nh,hs) and transposed -- consumed in place through their strides."""
    from flash_attention_ipu_b200._impl import ops
    B, T, nh, hs = 2, 256, 6, 64
    torch.manual_seed(SEED)
    c_attn = torch.randn(B, T, 3 * nh * hs, device="cuda", dtype=torch.bfloat16)
    q, k, v = (t.view(B, T, nh, hs).transpose(1, 2) for t in c_attn.split(nh * hs, dim=2))
    out, lse = ops.attention_fwd(q, k, v, hs ** -0.5, True, 128, 128)
    ref = ao.sdpa_reference(q.float().cpu(), k.float().cpu(), v.float().cpu())
    assert (out.float().cpu() - ref).abs().max().item() <= 1e-2
    # the output keeps q's (B, T, nh, hs)-transposed dense layout (what `y.transpose(1, 2).contiguous()`
    # in nanoGPT then turns into a free view): strides of a (B, T, nh, hs) tensor seen as (B, nh, T, hs)
    assert out.stride() == (T * nh * hs, hs, nh * hs, 1)


def test_memory_is_linear_in_sequence_length():
    """The reference's reason to exist (README.md:59-73; the OOM test at
    tests/python/test_scaled_dot_product_attention.py:116-153): no L x L tensor is materialised."""
    import flash_attention_ipu as fa
    G, L, D = 16, 16384, 64
    qkv = torch.randn(3, G, L, D, device="cuda", dtype=torch.bfloat16).requires_grad_(True)
    torch.cuda.synchronize(); torch.cuda.reset_peak_memory_stats()
    base = torch.cuda.memory_allocated()
    out = fa.flash_attention_qkv_packed(qkv, 16, 16)
    out.sum().backward()
    torch.cuda.synchronize()
    extra = torch.cuda.max_memory_allocated() - base
    scores_bytes = G * L * L * 2
    assert extra < 8 * qkv.numel() * 2, extra       # O(L): a few copies of qkv (+ fp32 dQ scratch)
    assert extra < scores_bytes / 20


def test_merge_kernel():
    native = _native_loaded()
    torch.manual_seed(3)
    o_a = torch.randn(4, 100, 64, device="cuda").bfloat16(); o_b = torch.randn_like(o_a)
    lse_a = torch.randn(4, 100, device="cuda"); lse_b = torch.randn(4, 100, device="cuda")
    lse_b[0, :10] = float("-inf")
    ro, rl = ao.merge_partials(o_a.float().cpu(), lse_a.cpu(), o_b.float().cpu(), lse_b.cpu())
    native.merge(o_a, lse_a, o_b, lse_b)
    assert (o_a.float().cpu() - ro).abs().max().item() < 2e-2
    assert (lse_a.cpu() - rl).abs().max().item() < 1e-5


def test_errors_are_loud():
    native = _native_loaded()
    q = torch.zeros(1, 1, 128, 512, device="cuda", dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="fa_b200_fwd failed"):
        native.fwd(native.make_problem(q, q, True, 1.0), q, q, q, q, torch.zeros(1, 1, 128, device="cuda"))


def test_merge_acc_kernel():
    """fa_b200_merge_acc (fp32 ring accumulators, sub-range of rows) vs the oracle's merge."""
    native = _native_loaded()
    torch.manual_seed(SEED)
    G, La, Lb, D, row0 = 3, 96, 40, 128, 56
    o_acc, lse_acc = torch.randn(G, La, D), torch.randn(G, La)
    lse_acc[0, 60] = float("-inf")  # empty accumulator row: takes the partial as is
    o_b, lse_b = torch.randn(G, Lb, D).bfloat16(), torch.randn(G, Lb)
    lse_b[1, 3] = float("-inf")     # empty partial row: accumulator unchanged
    o_ref, lse_ref = o_acc.clone(), lse_acc.clone()
    o_m, lse_m = ao.merge_partials(o_acc[:, row0:row0 + Lb], lse_acc[:, row0:row0 + Lb], o_b.float(), lse_b)
    o_ref[:, row0:row0 + Lb], lse_ref[:, row0:row0 + Lb] = o_m, lse_m
    o_d, lse_d = o_acc.cuda(), lse_acc.cuda()
    native.merge_acc(o_d, lse_d, o_b.cuda(), lse_b.cuda(), row0)
    torch.cuda.synchronize()
    assert torch.allclose(lse_d.cpu(), lse_ref, atol=1e-5)
    ok = ~torch.isnan(o_ref)
    assert torch.allclose(o_d.cpu()[ok], o_ref[ok], atol=1e-5)
    assert not torch.isnan(o_d).any()


@pytest.mark.parametrize("world,T,D", [(2, 1024, 128), (4, 2048, 64), (8, 4096, 128)])
def test_emulated_ring_vs_single_gpu_and_oracle(world, T, D):
    """Config-5 schedule (zig-zag shards, one block per hop, fp32 LSE merge, travelling dK/dV) with
    every rank stepped on one GPU: must agree with the un-sharded kernel and with the fp32 oracle."""
    _native_loaded()
    from flash_attention_ipu_b200 import ring
    import flash_attention_ipu.auto  # noqa: F401  (SDPA override = the un-sharded CUDA path)
    torch.manual_seed(SEED)
    B, H = 1, 2
    q, k, v, dout = (torch.randn(B, H, T, D).bfloat16() for _ in range(4))
    sh = lambda t: [ring.zigzag_shard(t.cuda(), r, world) for r in range(world)]
    outs, grads = ring.ring_attention_emulated(sh(q), sh(k), sh(v), sh(dout))
    torch.cuda.synchronize()
    out = ring.zigzag_unshard(outs).float().cpu()
    dq, dk, dv = (ring.zigzag_unshard([g[i] for g in grads]).float().cpu() for i in range(3))
    # fp32 oracle on the same rounded inputs
    qf, kf, vf = (t.float().requires_grad_(True) for t in (q, k, v))
    ref = ao.sdpa_reference(qf, kf, vf, is_causal=True)
    ref.backward(dout.float())
    assert (out - ref.detach()).abs().max().item() <= 1e-2
    for got, want in ((dq, qf.grad), (dk, kf.grad), (dv, vf.grad)):
        assert rel_fro(got, want) <= 1e-2
    # un-sharded CUDA kernel
    qd, kd, vd = (t.cuda().requires_grad_(True) for t in (q, k, v))
    o1 = torch.nn.functional.scaled_dot_product_attention(qd, kd, vd, is_causal=True)
    o1.backward(dout.cuda())
    assert (out - o1.detach().float().cpu()).abs().max().item() <= 1e-2
    for got, want in ((dq, qd.grad), (dk, kd.grad), (dv, vd.grad)):
        assert rel_fro(got, want.float().cpu()) <= 1e-2


def test_nanogpt_demo_loss_curve_parity():
    """SURVEY.md 8f-2 / notebook Appendix A of the reference: training the nanoGPT-shaped model with
    the SDPA override follows the un-patched loss curve (same seed, bf16 autocast), the loss goes
    down, and the attention launches really are ours."""
    import importlib.util
    import pathlib
    spec = importlib.util.spec_from_file_location(
        "train_nanogpt_b200", pathlib.Path(__file__).resolve().parent.parent / "demo" / "train_nanogpt_b200.py")
    demo = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(demo)
    kw = dict(steps=40, n_layer=2, batch=16, block=256, n_head=6, n_embd=384)
    base, _, l0, _ = demo.train(False, **kw)
    ours, _, l1, _ = demo.train(True, **kw)
    assert l0 == 0 and l1 >= 2 * 2 * 40          # fwd + bwd (+ prep/convert) per layer per step
    assert ours[-1] < ours[0] - 0.5               # it learns
    assert max(abs(a - b) for a, b in zip(base[:10], ours[:10])) <= 0.1    # same curve (bf16 training noise is ~0.04)
    assert abs(sum(base[-10:]) - sum(ours[-10:])) / 10 <= 0.05


def test_host_streamed_attention_matches_device_path():
    """flash_attention_ipu_b200/streaming.py (what bench.py's e2e leg calls): chunked, overlapped
    host->device->host pipeline gives exactly the result of the one-shot device call."""
    import flash_attention_ipu as fa
    from flash_attention_ipu_b200.streaming import HostStreamedAttention
    torch.manual_seed(SEED)
    G, L, D = 12, 384, 128
    h_qkv = torch.randn(3, G, L, D).bfloat16().pin_memory()
    h_do = torch.randn(G, L, D).bfloat16().pin_memory()
    h_out = torch.empty(G, L, D, dtype=torch.bfloat16).pin_memory()
    h_dqkv = torch.empty(3, G, L, D, dtype=torch.bfloat16).pin_memory()
    pipe = HostStreamedAttention(G, L, D, groups_per_chunk=4)
    for _ in range(2):  # second run re-uses the staging buffers
        pipe.run(h_qkv, h_do, h_out, h_dqkv)
    torch.cuda.synchronize()
    # the ramped schedule (1, 1, 2 | 4 | 2, 1, 1 groups: shorter fill and drain) gives the same bytes
    h_out2, h_dqkv2 = torch.empty_like(h_out).pin_memory(), torch.empty_like(h_dqkv).pin_memory()
    ramped = HostStreamedAttention(G, L, D, groups_per_chunk=4, ramp=True)
    assert [n for _, n in ramped.chunks] == [1, 1, 2, 4, 2, 1, 1]
    ramped.run(h_qkv, h_do, h_out2, h_dqkv2)
    assert torch.equal(h_out2, h_out) and torch.equal(h_dqkv2[1:], h_dqkv[1:])
    assert (h_dqkv2[0].float() - h_dqkv[0].float()).abs().max().item() <= 2e-2
    x = h_qkv.cuda().requires_grad_(True)
    out = fa.flash_attention_qkv_packed(x, 1, 1)
    out.backward(h_do.cuda())
    torch.cuda.synchronize()
    assert torch.equal(h_out, out.detach().cpu())
    # dQ accumulates by fp32 atomics in a different order per launch: equal up to that
    assert (h_dqkv.float() - x.grad.float().cpu()).abs().max().item() <= 2e-2
    assert torch.equal(h_dqkv[1:], x.grad[1:].cpu())


@pytest.mark.parametrize("Hq,Hkv,L,D", [(8, 2, 512, 128), (6, 1, 384, 64), (4, 4, 256, 128)])
@pytest.mark.parametrize("causal", [True, False])
def test_grouped_query_attention(Hq, Hkv, L, D, causal):
    """Grouped-/multi-query attention (fewer K/V heads) through the SDPA override with
    enable_gqa=True -- an extension: the reference raises for q.shape != k.shape
    (_impl/scaled_dot_product_attention.py:33-37).  Oracle: the same math with K/V heads repeated."""
    native = _native_loaded()
    from flash_attention_ipu_b200._impl import scaled_dot_product_attention as sd
    sd.install()
    launches0 = native.launch_count()
    torch.manual_seed(SEED)
    B = 2
    q = torch.randn(B, Hq, L, D).bfloat16()
    k, v = torch.randn(B, Hkv, L, D).bfloat16(), torch.randn(B, Hkv, L, D).bfloat16()
    dout = torch.randn(B, Hq, L, D).bfloat16()
    qd, kd, vd = (t.cuda().requires_grad_(True) for t in (q, k, v))
    out = torch.nn.functional.scaled_dot_product_attention(qd, kd, vd, is_causal=causal, enable_gqa=True)
    out.backward(dout.cuda())
    torch.cuda.synchronize()
    assert native.launch_count() - launches0 >= 2, "the override did not route to the CUDA kernels"
    qf, kf, vf = (t.float().requires_grad_(True) for t in (q, k, v))
    rep = Hq // Hkv
    ref = ao.sdpa_reference(qf, kf.repeat_interleave(rep, dim=1), vf.repeat_interleave(rep, dim=1), is_causal=causal)
    ref.backward(dout.float())
    assert (out.float().cpu() - ref.detach()).abs().max().item() <= 1e-2
    for got, want in ((qd.grad, qf.grad), (kd.grad, kf.grad), (vd.grad, vf.grad)):
        assert got.shape == want.shape
        assert rel_fro(got.float().cpu(), want) <= 1e-2
    with pytest.raises(NotImplementedError):  # without the flag the reference behaviour is kept
        torch.nn.functional.scaled_dot_product_attention(qd, kd[:, :1], vd[:, :1], is_causal=causal)


def _blocked_reference(q, k, v, dout, causal, shift, scale):
    """fp64 attention of one (query block, key block) pair: key c visible to query r iff
    c <= r + shift; a row that sees nothing gives out = 0, lse = -inf and no gradient."""
    q, k, v = (t.double().requires_grad_(True) for t in (q, k, v))
    s = torch.einsum("bhqd,bhkd->bhqk", q, k) * scale
    if causal:
        r = torch.arange(q.shape[2])[:, None]
        c = torch.arange(k.shape[2])[None, :]
        s = s.masked_fill(c > r + shift, float("-inf"))
    lse = torch.logsumexp(s, dim=-1)
    p = torch.where(torch.isinf(lse)[..., None], torch.zeros_like(s), torch.exp(s - lse[..., None].clamp_min(-1e300)))
    out = torch.einsum("bhqk,bhkd->bhqd", p, v)
    out.backward(dout.double())
    return out.detach(), lse.detach(), q.grad, k.grad, v.grad


@pytest.mark.parametrize("B,H,Lq,Lk,D,causal,q0,k0,dtype", [
    (2, 3, 200, 200, 128, True, 0, 0, "fp32"),
    (1, 2, 331, 331, 64, True, 0, 0, "fp32"),
    (1, 2, 70, 150, 40, False, 0, 0, "fp32"),
    (1, 2, 100, 130, 96, True, 64, 0, "fp32"),     # query block later than the key block
    (1, 1, 90, 130, 24, True, 0, 40, "fp32"),      # first 40 rows see no key at all
    (1, 1, 1, 1, 8, True, 0, 0, "fp32"),
    (2, 2, 65, 65, 72, True, 0, 0, "bf16"),        # 16-bit input, head dim off the tensor path
    (1, 2, 129, 257, 100, False, 0, 0, "fp16"),
])
def test_tiled_simt_kernels(B, H, Lq, Lk, D, causal, q0, k0, dtype):
    """csrc/fa_tiled_simt.cuh through the C ABI: ragged lengths, padded head dims, shifted causal
    diagonals (the ring's block positions) and rows with nothing visible, against fp64 math."""
    native = _native_loaded()
    torch.manual_seed(SEED + Lq + D)
    dt = DT[dtype]
    q, k, v, dout = (torch.randn(B, H, L, D).to(dt) for L in (Lq, Lk, Lk, Lq))
    scale = D ** -0.5
    ref = _blocked_reference(q.float(), k.float(), v.float(), dout.float(), causal, q0 - k0, scale)
    qd, kd, vd, dd = (t.cuda() for t in (q, k, v, dout))
    prob = native.make_problem(qd, kd, causal, scale, 0, 0, q0, k0)
    assert native.kernel_family(prob) == 0
    o = torch.empty_like(qd)
    lse = torch.empty(B, H, Lq, device="cuda", dtype=torch.float32)
    native.fwd(prob, qd, kd, vd, o, lse)
    dq, dk, dv = torch.empty_like(qd), torch.empty_like(kd), torch.empty_like(vd)
    ws = torch.empty(max(native.bwd_workspace_bytes(prob), 4), device="cuda", dtype=torch.uint8)
    native.bwd(prob, qd, kd, vd, o, dd, lse, dq, dk, dv, ws)
    torch.cuda.synchronize()
    atol = 2e-4 if dtype == "fp32" else 3e-2
    assert torch.equal(torch.isinf(lse.cpu()), torch.isinf(ref[1]))
    fin = ~torch.isinf(ref[1])
    assert (lse.cpu().double()[fin] - ref[1][fin]).abs().max().item() <= (1e-4 if dtype == "fp32" else 1e-2)
    for got, want in zip((o, dq, dk, dv), (ref[0], ref[2], ref[3], ref[4])):
        assert (got.cpu().double() - want).abs().max().item() <= atol


def test_config1_full_size_fp32():
    """BASELINE config 1 at its full size (B=64, T=256, 6 heads, d=64, fp32, causal), as nanoGPT
    calls it (strided views of one c_attn output), against torch CPU SDPA in fp32."""
    import torch.nn.functional as F
    import flash_attention_ipu_b200._impl.scaled_dot_product_attention as mod
    native = _native_loaded()
    B, T, nh, hs = 64, 256, 6, 64
    torch.manual_seed(SEED)
    c_attn = torch.randn(B, T, 3 * nh * hs)
    dout = torch.randn(B, nh, T, hs)
    xr = c_attn.clone().requires_grad_(True)
    q, k, v = (t.view(B, T, nh, hs).transpose(1, 2) for t in xr.split(nh * hs, dim=2))
    ref = ao.sdpa_reference(q, k, v)
    (ref * dout).sum().backward()
    was = getattr(F.scaled_dot_product_attention, "__wrapped__", None) is not None
    mod.install()
    try:
        before = native.launch_count()
        x = c_attn.cuda().requires_grad_(True)
        q, k, v = (t.view(B, T, nh, hs).transpose(1, 2) for t in x.split(nh * hs, dim=2))
        out = F.scaled_dot_product_attention(q, k, v, is_causal=True)
        out.backward(dout.cuda())
        torch.cuda.synchronize()
        assert native.launch_count() - before == 3  # forward, dQ (+delta), dK/dV
    finally:
        if not was:
            mod.uninstall()
    assert (out.detach().cpu() - ref.detach()).abs().max().item() <= 2e-5
    assert (x.grad.cpu() - xr.grad).abs().max().item() <= 1e-4


@pytest.mark.gpu
def test_tensor_map_cache_reuses_descriptors():
    """The shim caches TMA descriptors by (pointer, dtype, shape, strides, box): a second call on the
    same buffers must not encode again, and must give the same result (threading row)."""
    import flash_attention_ipu as fa
    from flash_attention_ipu_b200._impl import native
    torch.manual_seed(7)
    qkv = torch.randn(3, 2, 256, 128, device="cuda", dtype=torch.bfloat16)
    out1 = fa.flash_attention_qkv_packed(qkv, 2, 2).clone()
    h0, m0 = native.tmap_cache_stats()
    out2 = fa.flash_attention_qkv_packed(qkv, 2, 2)
    h1, m1 = native.tmap_cache_stats()
    assert h1 - h0 >= 3, "q, k, v descriptors were re-encoded"
    assert torch.equal(out1, out2)


# ------------------------------------------------------------------ round 2: BASELINE.json shapes at their own sizes
def test_config3_shape_slice_out_and_grads():
    """BASELINE config 3's own length and head dim (T=8192, d=128, bf16, causal), G=2 of the 128 groups:
    out AND dq/dk/dv against the blocked restatement of the reference's chunked algorithm
    (cpp/flash_attention_qkv_packed.cpp:229-328, :410-497) with 8 x 8 chunks."""
    G, L, D = 2, 8192, 128
    qkv, grad = packed_inputs(G, L, D)
    qkv, grad = qkv.bfloat16().float(), grad.bfloat16().float()
    out, dqkv = run_packed(qkv, grad, torch.bfloat16, 8, 8)
    ref, _ = ao.chunked_attention_forward(qkv, 8, 8)
    dref = ao.chunked_attention_backward(grad, qkv, 8, 8)
    assert (out - ref).abs().max().item() <= 1e-2
    for t in range(3):
        assert rel_fro(dqkv[t], dref[t]) < 1e-2
        assert (dqkv[t] - dref[t]).abs().max().item() <= 1.0  # the reference's own 16-bit gradient bound


@pytest.mark.parametrize("name,G,L,D,causal", [("config2_gpt2_small", 8, 1024, 64, True),
                                               ("config4_bert_large", 8, 512, 64, False)])
@pytest.mark.parametrize("api", ["packed", "sdpa"])
def test_config2_and_config4_exact_shapes(name, G, L, D, causal, api):
    """BASELINE configs 2 (B=8, T=1024, 12 heads, d=64, causal) and 4 (B=32, T=512, 16 heads, d=64,
    NON-causal: parity unpinned, the reference refuses it) at their own T / d, 8 of their groups,
    bf16, through both entry points."""
    import torch.nn.functional as F
    import flash_attention_ipu_b200._impl.scaled_dot_product_attention as mod
    qkv, grad = packed_inputs(G, L, D)
    qkv, grad = qkv.bfloat16().float(), grad.bfloat16().float()
    ref = ao.packed_attention_reference(qkv, causal)
    dref = ao.vanilla_attention_grad(grad, qkv, causal)
    if api == "packed":
        out, dqkv = run_packed(qkv, grad, torch.bfloat16, 2, 2, causal=causal)
    else:
        mod.install()
        try:
            x = qkv.view(3, 2, G // 2, L, D).to("cuda", torch.bfloat16).requires_grad_(True)
            o = F.scaled_dot_product_attention(x[0], x[1], x[2], is_causal=causal, scale=1.0)
            o.backward(grad.view(2, G // 2, L, D).to("cuda", torch.bfloat16))
            torch.cuda.synchronize()
        finally:
            mod.uninstall()
        out, dqkv = o.detach().float().cpu().view(G, L, D), x.grad.float().cpu().view(3, G, L, D)
    assert (out - ref).abs().max().item() <= 1e-2
    for t in range(3):
        assert rel_fro(dqkv[t], dref[t]) < 1e-2


def test_c_abi_tensor_family_direct():
    """fa_b200_fwd / fa_b200_bwd called through ctypes with raw pointers (no torch op in between)
    on a problem that dispatches to the tcgen05 family, ragged lengths and a shifted diagonal."""
    native = _native_loaded()
    torch.manual_seed(SEED)
    B, H, Lq, Lk, D, q0, k0 = 1, 3, 200, 328, 128, 128, 0
    q, k, v, dout = (torch.randn(B, H, L, D).bfloat16() for L in (Lq, Lk, Lk, Lq))
    scale = D ** -0.5
    ref = _blocked_reference(q.float(), k.float(), v.float(), dout.float(), True, q0 - k0, scale)
    qd, kd, vd, dd = (t.cuda() for t in (q, k, v, dout))
    prob = native.make_problem(qd, kd, True, scale, 0, 0, q0, k0)
    assert native.kernel_family(prob) == 1
    o = torch.empty_like(qd)
    lse = torch.empty(B, H, Lq, device="cuda", dtype=torch.float32)
    native.fwd(prob, qd, kd, vd, o, lse)
    dq, dk, dv = torch.empty_like(qd), torch.empty_like(kd), torch.empty_like(vd)
    ws = torch.empty(native.bwd_workspace_bytes(prob), device="cuda", dtype=torch.uint8)
    native.bwd(prob, qd, kd, vd, o, dd, lse, dq, dk, dv, ws)
    torch.cuda.synchronize()
    assert (lse.cpu().double() - ref[1]).abs().max().item() <= 1e-2
    assert (o.cpu().double() - ref[0]).abs().max().item() <= 1e-2
    for got, want in zip((dq, dk, dv), (ref[2], ref[3], ref[4])):
        assert rel_fro(got.cpu().double(), want) < 1e-2
    cyc, ns = native.bwd_sm_clock(prob, ws)  # the measurement hook reports a plausible SM clock
    assert cyc > 0 and ns > 0 and 300 < cyc / ns * 1e3 < 3000


def test_c_abi_test_binary():
    """build/tests (tests/cuda/test_flash_attention.cu: the reference's C++ case matrix and
    thresholds, tests/cpp/test_flash_attention.cpp:243-313, through the C ABI) run as a subprocess."""
    import pathlib
    import subprocess
    exe = pathlib.Path(__file__).resolve().parent.parent / "build" / "tests"
    assert exe.exists(), "run `make`"
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "0 failed" in r.stdout, r.stdout[-2000:]


def test_strongly_negative_scores_ragged_noncausal():
    """rows whose LSE is far below zero (q = -10 k) with a ragged, non-causal Lk: the
    zero-filled key rows of the last KV tile must not leak exp(-lse) into dQ (NaN / inf)."""
    G, L, D = 2, 200, 128
    torch.manual_seed(SEED)
    k = torch.randn(G, L, D)
    qkv = torch.stack([-10.0 * k * D ** -0.5, k, torch.randn(G, L, D)]).bfloat16().float()
    grad = torch.randn(G, L, D).bfloat16().float()
    for causal in (False, True):
        out, dqkv = run_packed(qkv, grad, torch.bfloat16, causal=causal)
        assert torch.isfinite(out).all() and torch.isfinite(dqkv).all()
        ref = ao.packed_attention_reference(qkv, causal)
        dref = ao.vanilla_attention_grad(grad, qkv, causal)
        assert (out - ref).abs().max().item() <= 2e-2
        assert rel_fro(dqkv, dref) < 2e-2


def test_sdpa_expanded_and_size1_strides():
    """K/V expanded along heads (the old multi-query idiom, stride 0) and operands
    with size-1 dims of arbitrary stride go through the override like stock torch handles them."""
    import torch.nn.functional as F
    import flash_attention_ipu_b200._impl.scaled_dot_product_attention as mod
    torch.manual_seed(SEED)
    B, H, L, D = 2, 4, 256, 64
    q = torch.randn(B, H, L, D, device="cuda", dtype=torch.bfloat16)
    k1 = torch.randn(B, 1, L, D, device="cuda", dtype=torch.bfloat16)
    v1 = torch.randn(B, 1, L, D, device="cuda", dtype=torch.bfloat16)
    k, v = k1.expand(B, H, L, D), v1.expand(B, H, L, D)
    ref = ao.sdpa_reference(q.float().cpu(), k.float().cpu(), v.float().cpu())
    mod.install()
    try:
        out = F.scaled_dot_product_attention(q, k, v, is_causal=True)
        # batch of 1 whose stride is not a multiple of 16 bytes (a slice of a padded buffer)
        buf = torch.randn(1, H * L * D + 3, device="cuda", dtype=torch.bfloat16)
        q1 = buf.as_strided((1, H, L, D), (H * L * D + 3, L * D, D, 1))
        out1 = F.scaled_dot_product_attention(q1, q1, q1, is_causal=True)
        ref1 = ao.sdpa_reference(q1.float().cpu(), q1.float().cpu(), q1.float().cpu())
        # a non-positive scale is torch's business, not ours: routed to the original function
        from flash_attention_ipu_b200._impl import native
        before = native.launch_count()
        F.scaled_dot_product_attention(q, k, v, is_causal=True, scale=-0.1)
        routed_to_torch = native.launch_count() == before
    finally:
        mod.uninstall()
    assert (out.float().cpu() - ref).abs().max().item() <= 1e-2
    # q = k = v of unscaled N(0,1): the softmax is sharply peaked and |out| reaches 3-4, where one bf16
    # ulp is 1.6e-2 -- a relative bound is the meaningful one here
    assert rel_fro(out1.float().cpu(), ref1) < 1e-2
    assert routed_to_torch


def test_ops_check_their_operands():
    """the public ops refuse operands that do not describe one problem instead of
    reading the wrong rows."""
    from flash_attention_ipu_b200._impl import ops
    q = torch.randn(1, 2, 128, 64, device="cuda", dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="share a dtype"):
        ops.attention_fwd(q, q.half(), q, 1.0, True, 0, 0)
    with pytest.raises(RuntimeError, match="does not match"):
        ops.attention_fwd(q, q, q[:, :, :64], 1.0, True, 0, 0)
    out, lse = ops.attention_fwd(q, q, q, 0.125, True, 0, 0)
    with pytest.raises(RuntimeError, match="lse"):
        ops.attention_bwd(out, q, q, q, out, lse[:, :, :64], 0.125, True, 0, 0)
    with pytest.raises(RuntimeError, match="out / dout"):
        ops.attention_bwd(out[:, :, :64], q, q, q, out, lse, 0.125, True, 0, 0)
    # a non-contiguous lse (a slice along L) is accepted and read correctly
    big, lse_big = ops.attention_fwd(torch.cat([q, q], 2), torch.cat([q, q], 2), torch.cat([q, q], 2), 0.125, True, 0, 0)
    g1 = ops.attention_bwd(out, q, q, q, big[:, :, :128], lse_big[:, :, :128], 0.125, True, 0, 0)
    g2 = ops.attention_bwd(out, q, q, q, out, lse, 0.125, True, 0, 0)
    for a, b in zip(g1, g2):
        assert rel_fro(a.float(), b.float()) < 1e-5


def test_bwd_dq_accum_matches_bwd():
    """fa_b200_bwd_dq_accum: dQ reduce-added in place into a caller-owned fp32 accumulator (unscaled),
    twice, equals 2 x the ordinary backward's dQ; dK / dV identical to the ordinary call."""
    native = _native_loaded()
    torch.manual_seed(SEED)
    B, H, L, D = 1, 3, 328, 128
    q, k, v, dout = (torch.randn(B, H, L, D, device="cuda").bfloat16() for _ in range(4))
    scale = D ** -0.5
    prob = native.make_problem(q, k, True, scale)
    o = torch.empty_like(q)
    lse = torch.empty(B, H, L, device="cuda", dtype=torch.float32)
    native.fwd(prob, q, k, v, o, lse)
    ws = torch.empty(native.bwd_workspace_bytes(prob), device="cuda", dtype=torch.uint8)
    dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
    native.bwd(prob, q, k, v, o, dout, lse, dq, dk, dv, ws)
    # accumulator = a strided slice of a larger fp32 buffer, pre-filled: the kernel must add, not overwrite
    big = torch.full((B, H, 2 * L, D), 0.5, device="cuda", dtype=torch.float32)
    acc = big[:, :, L:]
    dk2, dv2 = torch.empty_like(k), torch.empty_like(v)
    native.bwd_dq_accum(prob, q, k, v, o, dout, lse, acc, dk2, dv2, ws)
    native.bwd_dq_accum(prob, q, k, v, o, dout, lse, acc, dk2, dv2, ws)
    torch.cuda.synchronize()
    assert torch.equal(big[:, :, :L], torch.full_like(big[:, :, :L], 0.5))
    got = (acc - 0.5) * scale / 2
    assert rel_fro(got.cpu(), dq.float().cpu()) < 4e-3   # dq is the bf16 rounding of the same sum
    assert torch.equal(dk2, dk) and torch.equal(dv2, dv)


@pytest.mark.parametrize("G,L,D,causal", [(8, 4096, 128, True), (3, 2048, 64, True), (2, 1000, 128, False)])
def test_deterministic_dq_is_bit_reproducible(G, L, D, causal):
    """fa_b200_set_deterministic: dQ partials added in a fixed order behind per-tile semaphores --
    repeated runs are bit-identical (the default arrival-order sum is only reproducible to ~1e-7),
    and the result agrees with the default mode to fp32 summation noise."""
    import flash_attention_ipu_b200 as pkg
    qkv, grad = packed_inputs(G, L, D)
    base = run_packed(qkv, grad, torch.bfloat16, causal=causal)
    prev = pkg.set_deterministic(True)
    try:
        runs = [run_packed(qkv, grad, torch.bfloat16, causal=causal) for _ in range(4)]
    finally:
        pkg.set_deterministic(prev)
    for out, dqkv in runs[1:]:
        assert torch.equal(out, runs[0][0]) and torch.equal(dqkv, runs[0][1])
    assert torch.equal(runs[0][0], base[0])
    assert rel_fro(runs[0][1], base[1]) < 1e-3  # same sum, different order, then one bf16 rounding
    # torch's own switch selects the same mode for the torch ops
    torch.use_deterministic_algorithms(True, warn_only=True)
    try:
        again = run_packed(qkv, grad, torch.bfloat16, causal=causal)
    finally:
        torch.use_deterministic_algorithms(False)
    assert torch.equal(again[1], runs[0][1])


def _masked_reference(q, k, v, dout, lens, causal, scale):
    """fp64 attention with a key-padding mask: keys >= lens[b] do not exist."""
    q, k, v = (t.double().requires_grad_(True) for t in (q, k, v))
    s = torch.einsum("bhqd,bhkd->bhqk", q, k) * scale
    Lq, Lk = q.shape[2], k.shape[2]
    pad = torch.arange(Lk)[None, None, None, :] >= lens.view(-1, 1, 1, 1)
    s = s.masked_fill(pad, float("-inf"))
    if causal:
        s = s.masked_fill(torch.arange(Lk)[None, :] > torch.arange(Lq)[:, None], float("-inf"))
    out = torch.einsum("bhqk,bhkd->bhqd", torch.softmax(s, dim=-1), v)
    out.backward(dout.double())
    return out.detach(), q.grad, k.grad, v.grad


@pytest.mark.parametrize("B,H,L,D,causal,lens", [
    (3, 2, 512, 128, False, [512, 200, 129]),   # BERT-style padding, ragged inside a tile and on a tile edge
    (2, 4, 384, 64, False, [384, 1]),
    (2, 2, 512, 128, True, [300, 512]),         # padding on top of the causal mask
])
def test_key_padding_mask_via_sdpa(B, H, L, D, causal, lens):
    """attn_mask as a boolean key-padding mask (the reference refuses every mask,
    _impl/scaled_dot_product_attention.py:21-24): out and all gradients vs fp64 math; padded keys get
    dk = dv = 0.  (Other mask forms: test_attn_mask_through_sdpa_override.)"""
    import torch.nn.functional as F
    import flash_attention_ipu_b200._impl.scaled_dot_product_attention as mod
    native = _native_loaded()
    torch.manual_seed(SEED + L)
    q, k, v, dout = (torch.randn(B, H, L, D).bfloat16() for _ in range(4))
    lens_t = torch.tensor(lens)
    ref = _masked_reference(q.float(), k.float(), v.float(), dout.float(), lens_t, causal, D ** -0.5)
    mask = (torch.arange(L)[None, :] < lens_t[:, None]).view(B, 1, 1, L).cuda()
    qd, kd, vd = (t.cuda().requires_grad_(True) for t in (q, k, v))
    mod.install()
    try:
        before = native.launch_count()
        out = F.scaled_dot_product_attention(qd, kd, vd, attn_mask=mask, is_causal=causal)
        out.backward(dout.cuda())
        torch.cuda.synchronize()
        assert native.launch_count() > before
        with pytest.raises(NotImplementedError, match="attn_mask"):   # a mask that does not broadcast to (B, H, Lq, Lk)
            F.scaled_dot_product_attention(qd, kd, vd, attn_mask=torch.zeros(L, L // 2, device="cuda", dtype=torch.bfloat16))
    finally:
        mod.uninstall()
    # causal rows whose visible keys are all real: every row has at least key 0 (lens >= 1)
    assert (out.detach().float().cpu().double() - ref[0]).abs().max().item() <= 1e-2
    for got, want in zip((qd.grad, kd.grad, vd.grad), ref[1:]):
        assert rel_fro(got.float().cpu().double(), want) < 1e-2
    for b, n in enumerate(lens):
        assert torch.count_nonzero(kd.grad[b, :, n:]) == 0 and torch.count_nonzero(vd.grad[b, :, n:]) == 0


def test_kv_lens_packed_api_and_skipped_tiles():
    """kv_lens through the packed API: same result as physically truncating K / V, for every group."""
    import flash_attention_ipu as fa
    torch.manual_seed(SEED)
    G, L, D = 3, 640, 128
    qkv = torch.randn(3, G, L, D, device="cuda", dtype=torch.bfloat16)
    qkv[0] *= D ** -0.5
    lens = torch.tensor([640, 257, 128], device="cuda", dtype=torch.int32)
    x = qkv.clone().requires_grad_(True)
    out = fa.flash_attention_qkv_packed(x, 1, 1, causal=False, kv_lens=lens)
    g = torch.randn_like(out)
    out.backward(g)
    from flash_attention_ipu_b200._impl import ops
    for n in range(G):
        m = int(lens[n])
        q1, k1, v1 = (qkv[i, n][None, None].clone().requires_grad_(True) for i in range(3))
        o1, _ = ops.attention_fwd(q1, k1[:, :, :m], v1[:, :, :m], 1.0, False, 0, 0)
        o1.backward(g[n][None, None])
        assert torch.equal(o1[0, 0], out[n])
        # dQ: fp32 partials added in arrival order, then rounded to bf16 -- single-ulp flips are legitimate
        assert rel_fro(x.grad[0, n].float(), q1.grad[0, 0].float()) < 1e-3
        assert torch.equal(x.grad[1, n, :m], k1.grad[0, 0, :m]) and torch.equal(x.grad[2, n, :m], v1.grad[0, 0, :m])
        assert torch.count_nonzero(x.grad[1:, n, m:]) == 0


def _dropout_reference(q, k, v, dout, keep, drop_scale, causal, shift, scale):
    """fp64 attention with dropout on the probabilities: out = drop_scale * (keep o softmax(S)) V; the
    LSE is that of the undropped scores.  `keep` is (B, H, Lq, Lk) bool; k, v may have fewer heads."""
    q, k, v = (t.double().requires_grad_(True) for t in (q, k, v))
    rep = q.shape[1] // k.shape[1]
    kk, vv = k.repeat_interleave(rep, dim=1), v.repeat_interleave(rep, dim=1)
    s = torch.einsum("bhqd,bhkd->bhqk", q, kk) * scale
    if causal:
        r = torch.arange(q.shape[2])[:, None]
        c = torch.arange(k.shape[2])[None, :]
        s = s.masked_fill(c > r + shift, float("-inf"))
    lse = torch.logsumexp(s, dim=-1)
    p = torch.where(torch.isinf(lse)[..., None], torch.zeros_like(s), torch.exp(s - lse[..., None].clamp_min(-1e300)))
    out = torch.einsum("bhqk,bhkd->bhqd", p * keep.double() * drop_scale, vv)
    out.backward(dout.double())
    return out.detach(), lse.detach(), q.grad, k.grad, v.grad


@pytest.mark.parametrize("B,H,Hkv,Lq,Lk,D,causal,q0,k0,dtype,p_drop", [
    (1, 3, 3, 200, 328, 128, True, 128, 0, "bf16", 0.25),   # ragged, shifted diagonal (a ring hop's positions)
    (2, 2, 2, 384, 384, 64, False, 0, 0, "fp16", 0.1),
    (1, 4, 2, 256, 256, 128, True, 0, 64, "bf16", 0.5),     # grouped-query, key block ahead of the queries
])
def test_dropout_matches_the_hash_oracle(B, H, Hkv, Lq, Lk, D, causal, q0, k0, dtype, p_drop):
    """fa_b200_problem.dropout_p / dropout_seed through the C ABI: the kernels' mask is the one
    oracle/attention_oracle.py:dropout_keep_mask restates (forward AND backward regenerate it), the LSE
    is that of the undropped scores, and out / dq / dk / dv match fp64 attention with that mask."""
    native = _native_loaded()
    torch.manual_seed(SEED + Lq + D)
    dt = DT[dtype]
    seed = 0x1234_5678_9ABC_DEF1
    q, dout = (torch.randn(B, H, Lq, D).to(dt) for _ in range(2))
    k, v = (torch.randn(B, Hkv, Lk, D).to(dt) for _ in range(2))
    scale = D ** -0.5
    keep, drop_scale = ao.dropout_keep_mask(seed, B * H, Lq, Lk, p_drop, q0, k0)
    keep = keep.view(B, H, Lq, Lk)
    assert abs(keep.float().mean().item() - (1 - p_drop)) < 0.02
    ref = _dropout_reference(q.float(), k.float(), v.float(), dout.float(), keep, drop_scale, causal, q0 - k0, scale)
    qd, kd, vd, dd = (t.cuda() for t in (q, k, v, dout))
    prob = native.make_problem(qd, kd, causal, scale, 0, 0, q0, k0, dropout_p=p_drop, dropout_seed=seed)
    assert native.kernel_family(prob) == 1
    o = torch.empty_like(qd)
    lse = torch.empty(B, H, Lq, device="cuda", dtype=torch.float32)
    native.fwd(prob, qd, kd, vd, o, lse)
    dq, dk, dv = torch.empty_like(qd), torch.empty_like(kd), torch.empty_like(vd)
    ws = torch.empty(native.bwd_workspace_bytes(prob), device="cuda", dtype=torch.uint8)
    native.bwd(prob, qd, kd, vd, o, dd, lse, dq, dk, dv, ws)
    torch.cuda.synchronize()
    finite = torch.isfinite(ref[1])
    assert torch.equal(torch.isfinite(lse.cpu()), finite)
    assert (lse.cpu().double()[finite] - ref[1][finite]).abs().max().item() <= 1e-2
    # a wrong mask bit moves an output element by O(1); the 16-bit rounding of P moves it by ~1e-2
    assert (o.cpu().double() - ref[0]).abs().max().item() <= 3e-2
    assert rel_fro(o.cpu().double(), ref[0]) < 1e-2
    for got, want in zip((dq, dk, dv), (ref[2], ref[3], ref[4])):
        assert rel_fro(got.cpu().double(), want) < 1e-2
    # the seed matters, and dropout_p = 0 is the plain kernel
    prob2 = native.make_problem(qd, kd, causal, scale, 0, 0, q0, k0, dropout_p=p_drop, dropout_seed=seed + 1)
    o2 = torch.empty_like(qd)
    native.fwd(prob2, qd, kd, vd, o2, lse)
    torch.cuda.synchronize()
    assert not torch.equal(o2, o)
    # refused off the tensor path
    qf = torch.randn(1, 1, 64, 40, device="cuda")
    with pytest.raises(RuntimeError, match="dropout"):
        native.fwd(native.make_problem(qf, qf, True, 1.0, dropout_p=0.1), qf, qf, qf, torch.empty_like(qf),
                   torch.empty(1, 1, 64, device="cuda"))


def test_dropout_through_sdpa_override_and_packed_api():
    """dropout_p > 0 through the public entry points: F.scaled_dot_product_attention (grouped-query, seed
    drawn from torch's CPU generator, hence repeatable under torch.manual_seed) and
    flash_attention_qkv_packed(dropout_p=, dropout_seed=); autograd carries the seed to the backward."""
    import torch.nn.functional as F
    import flash_attention_ipu as fa
    import flash_attention_ipu_b200._impl.scaled_dot_product_attention as mod
    native = _native_loaded()
    torch.manual_seed(SEED)
    B, H, Hkv, L, D, p_drop = 2, 4, 2, 256, 64, 0.2
    q, dout = (torch.randn(B, H, L, D).bfloat16() for _ in range(2))
    k, v = (torch.randn(B, Hkv, L, D).bfloat16() for _ in range(2))
    mod.install()
    try:
        def run(seed_for_torch):
            torch.manual_seed(seed_for_torch)
            xs = [t.cuda().requires_grad_(True) for t in (q, k, v)]
            before = native.launch_count()
            out = F.scaled_dot_product_attention(*xs, dropout_p=p_drop, is_causal=True, enable_gqa=True)
            out.backward(dout.cuda())
            torch.cuda.synchronize()
            assert native.launch_count() - before >= 3  # forward, prep, backward (+ convert): our kernels, not torch's
            return [out.detach().cpu()] + [x.grad.cpu() for x in xs]
        a, b, c = run(11), run(11), run(12)
        assert torch.equal(a[0], b[0]) and torch.equal(a[2], b[2]) and torch.equal(a[3], b[3])  # same seed, same mask
        assert not torch.equal(a[0], c[0])
        torch.manual_seed(11)
        seed = int(torch.randint(0, 2 ** 62, (1,)).item())  # what the override drew
    finally:
        mod.uninstall()
    keep, drop_scale = ao.dropout_keep_mask(seed, B * H, L, L, p_drop)
    ref = _dropout_reference(q.float(), k.float(), v.float(), dout.float(), keep.view(B, H, L, L), drop_scale, True, 0, D ** -0.5)
    assert rel_fro(a[0].double(), ref[0]) < 1e-2
    for got, want in zip(a[1:], ref[2:]):
        assert rel_fro(got.double(), want) < 1e-2

    # packed API, explicit seed
    G, L, D = 3, 256, 128
    qkv = torch.randn(3, G, L, D)
    qkv[0] *= D ** -0.5  # the packed API does not scale (softmax_scale = 1): the caller pre-scales q
    qkv = qkv.bfloat16()
    g_out = torch.randn(G, L, D).bfloat16()
    out, dqkv = run_packed(qkv, g_out, torch.bfloat16, dropout_p=0.3, dropout_seed=99)
    keep, drop_scale = ao.dropout_keep_mask(99, G, L, L, 0.3)
    x = qkv.float()
    ref = _dropout_reference(x[0][None], x[1][None], x[2][None], g_out.float()[None], keep[None], drop_scale, True, 0, 1.0)
    assert rel_fro(out.double(), ref[0][0]) < 1e-2
    for t in range(3):
        assert rel_fro(dqkv[t].double(), ref[2 + t][0]) < 1e-2
    # fp32 / odd head dims keep the reference's refusal
    with pytest.raises(NotImplementedError, match="dropout_p"):
        mod._sdpa_impl(None, torch.randn(1, 2, 64, 32, device="cuda"), torch.randn(1, 2, 64, 32, device="cuda"),
                       torch.randn(1, 2, 64, 32, device="cuda"), None, 0.1, True)


def _bias_reference(q, k, v, dout, bias, causal, scale):
    """fp64 attention with an additive score bias (-inf = masked); k, v may have fewer heads."""
    q, k, v = (t.double().requires_grad_(True) for t in (q, k, v))
    rep = q.shape[1] // k.shape[1]
    kk, vv = k.repeat_interleave(rep, dim=1), v.repeat_interleave(rep, dim=1)
    s = torch.einsum("bhqd,bhkd->bhqk", q, kk) * scale + bias.double()
    if causal:
        r = torch.arange(q.shape[2])[:, None]
        c = torch.arange(k.shape[2])[None, :]
        s = s.masked_fill(c > r, float("-inf"))
    lse = torch.logsumexp(s, dim=-1)
    out = torch.einsum("bhqk,bhkd->bhqd", torch.exp(s - lse[..., None]), vv)
    out.backward(dout.double())
    return out.detach(), lse.detach(), q.grad, k.grad, v.grad


@pytest.mark.parametrize("B,H,Hkv,Lq,Lk,D,causal,dtype,bshape", [
    (2, 2, 2, 200, 328, 128, False, "bf16", "full"),    # ragged lengths, one bias row per (b, h, q)
    (1, 3, 3, 384, 384, 64, True, "fp16", "head"),      # (1, H, Lq, Lk): broadcast over the batch, with the causal mask
    (2, 4, 2, 256, 256, 128, False, "bf16", "keys"),    # grouped-query; (B, 1, 1, Lk): one row for every query (stride 0)
])
def test_additive_bias_through_the_c_abi(B, H, Hkv, Lq, Lk, D, causal, dtype, bshape):
    """fa_b200_problem.bias: S' = scale q.k + bias with -inf entries, forward and backward of both tcgen05
    kernels against fp64 math on the same (rounded) bias."""
    native = _native_loaded()
    torch.manual_seed(SEED + Lq + D)
    dt = DT[dtype]
    q, dout = (torch.randn(B, H, Lq, D).to(dt) for _ in range(2))
    k, v = (torch.randn(B, Hkv, Lk, D).to(dt) for _ in range(2))
    shape = {"full": (B, H, Lq, Lk), "head": (1, H, Lq, Lk), "keys": (B, 1, 1, Lk)}[bshape]
    bias = (2.0 * torch.randn(shape)).to(dt)
    hole = torch.rand(shape) < 0.3
    hole[..., 0] = False                                 # every row keeps key 0 (also under the causal mask)
    bias = bias.masked_fill(hole, float("-inf"))
    scale = D ** -0.5
    ref = _bias_reference(q.float(), k.float(), v.float(), dout.float(), bias.float().expand(B, H, Lq, Lk), causal, scale)
    qd, kd, vd, dd = (t.cuda() for t in (q, k, v, dout))
    bd = bias.cuda().expand(shape[0], shape[1], Lq, Lk)
    prob = native.make_problem(qd, kd, causal, scale, bias=bd)
    assert native.kernel_family(prob) == 1
    o = torch.empty_like(qd)
    lse = torch.empty(B, H, Lq, device="cuda", dtype=torch.float32)
    native.fwd(prob, qd, kd, vd, o, lse)
    dq, dk, dv = torch.empty_like(qd), torch.empty_like(kd), torch.empty_like(vd)
    ws = torch.empty(native.bwd_workspace_bytes(prob), device="cuda", dtype=torch.uint8)
    native.bwd(prob, qd, kd, vd, o, dd, lse, dq, dk, dv, ws)
    torch.cuda.synchronize()
    assert (lse.cpu().double() - ref[1]).abs().max().item() <= 1e-2
    assert (o.cpu().double() - ref[0]).abs().max().item() <= 2e-2
    assert rel_fro(o.cpu().double(), ref[0]) < 1e-2
    for got, want in zip((dq, dk, dv), (ref[2], ref[3], ref[4])):
        assert rel_fro(got.cpu().double(), want) < 1e-2
    # bias AND dropout (T5-style training): the same fp64 math with the oracle's keep mask on the probabilities
    seed, p_drop = 4242, 0.25
    keep, drop_scale = ao.dropout_keep_mask(seed, B * H, Lq, Lk, p_drop)
    x64 = [t.double().requires_grad_(True) for t in (q, k, v)]
    rep_kv = H // Hkv
    s64 = torch.einsum("bhqd,bhkd->bhqk", x64[0], x64[1].repeat_interleave(rep_kv, 1)) * scale + bias.double().expand(B, H, Lq, Lk)
    if causal:
        s64 = s64.masked_fill(torch.arange(Lk)[None, :] > torch.arange(Lq)[:, None], float("-inf"))
    ref_d = torch.einsum("bhqk,bhkd->bhqd", torch.softmax(s64, -1) * keep.view(B, H, Lq, Lk).double() * drop_scale,
                         x64[2].repeat_interleave(rep_kv, 1))
    ref_d.backward(dout.double())
    prob_d = native.make_problem(qd, kd, causal, scale, bias=bd, dropout_p=p_drop, dropout_seed=seed)
    native.fwd(prob_d, qd, kd, vd, o, lse)
    native.bwd(prob_d, qd, kd, vd, o, dd, lse, dq, dk, dv, ws)
    torch.cuda.synchronize()
    assert rel_fro(o.cpu().double(), ref_d.detach()) < 1e-2
    for got, want in zip((dq, dk, dv), x64):
        assert rel_fro(got.cpu().double(), want.grad) < 1e-2
    # refused: Lk not a multiple of 8, bias off the tensor path
    k7 = kd[:, :, :Lk - 1].contiguous()
    with pytest.raises(RuntimeError, match="bias"):
        bad = native.make_problem(qd, k7, causal, scale, bias=torch.zeros(B, H, Lq, Lk - 1, device="cuda", dtype=dt))
        native.fwd(bad, qd, k7, k7, o, lse)


def test_attn_mask_through_sdpa_override():
    """Float and boolean attn_mask (not of the key-padding form) through F.scaled_dot_product_attention: routed to
    the kernels (launch counter) and equal to torch's own masked SDPA run in fp32 on the CPU."""
    import torch.nn.functional as F
    import flash_attention_ipu_b200._impl.scaled_dot_product_attention as mod
    native = _native_loaded()
    torch.manual_seed(SEED)
    B, H, L, D = 2, 4, 256, 64
    q, k, v, dout = (torch.randn(B, H, L, D).bfloat16() for _ in range(4))
    idx = torch.arange(L)
    band = (idx[:, None] - idx[None, :]).abs() <= 96                      # (L, L) boolean sliding window
    alibi = -0.05 * (idx[:, None] - idx[None, :]).abs().float()           # float, one slope per head below
    masks = {
        "bool_2d": band,
        "bool_4d": band[None, None] & (torch.rand(B, 1, L, L) < 0.9) | torch.eye(L, dtype=torch.bool),
        "float_heads": (alibi[None, None] * torch.tensor([1.0, 2.0, 4.0, 8.0])[None, :, None, None]).bfloat16(),
    }
    orig = F.scaled_dot_product_attention.__wrapped__ if hasattr(F.scaled_dot_product_attention, "__wrapped__") \
        else F.scaled_dot_product_attention
    mod.install()
    try:
        for name, mask in masks.items():
            xs = [t.float().requires_grad_(True) for t in (q, k, v)]
            mref = mask if mask.dtype == torch.bool else mask.float()
            ref = orig(*xs, attn_mask=mref)
            ref.backward(dout.float())
            ys = [t.cuda().requires_grad_(True) for t in (q, k, v)]
            before = native.launch_count()
            out = F.scaled_dot_product_attention(*ys, attn_mask=mask.cuda())
            out.backward(dout.cuda())
            torch.cuda.synchronize()
            assert native.launch_count() - before >= 3, name
            assert (out.detach().float().cpu() - ref.detach()).abs().max().item() <= 2e-2, name
            for got, want in zip(ys, xs):
                assert rel_fro(got.grad.float().cpu(), want.grad) < 1e-2, name
        # key-padding shaped boolean masks with holes (left padding): the fast path off -> additive path
        mod.KEY_PADDING_FAST_PATH = False
        left = torch.ones(B, 1, 1, L, dtype=torch.bool)
        left[0, ..., :40] = False
        ys = [t.cuda() for t in (q, k, v)]
        out = F.scaled_dot_product_attention(*ys, attn_mask=left.cuda())
        ref = orig(q.float(), k.float(), v.float(), attn_mask=left)
        assert (out.float().cpu() - ref).abs().max().item() <= 2e-2
        # a mask that wants a gradient keeps the reference's refusal; a mask together with dropout runs on our kernels
        with pytest.raises(NotImplementedError):
            F.scaled_dot_product_attention(*ys, attn_mask=torch.zeros(L, L, device="cuda", dtype=torch.bfloat16, requires_grad=True))
        before = native.launch_count()
        both = F.scaled_dot_product_attention(*ys, attn_mask=band.cuda(), dropout_p=0.1)
        assert native.launch_count() > before and torch.isfinite(both.float()).all()
    finally:
        mod.KEY_PADDING_FAST_PATH = True
        mod.uninstall()


def test_item_order_chunks_do_not_change_a_group():
    """The persistent kernels walk the groups in L2-sized chunks (heaviest items of a chunk first; csrc/fa_api.cu:
    order_chunk).  40 groups of T=2048, d=128 are two chunks, the second one short: every group must come out
    exactly as when it is run alone (one chunk of one group) -- out, dK, dV bit-identical, dQ up to the
    arrival order of its fp32 partial sums."""
    import flash_attention_ipu as fa
    torch.manual_seed(SEED)
    G, L, D = 40, 2048, 128
    qkv = torch.randn(3, G, L, D, device="cuda", dtype=torch.bfloat16)
    qkv[0] *= D ** -0.5
    g_out = torch.randn(G, L, D, device="cuda", dtype=torch.bfloat16)
    x = qkv.clone().requires_grad_(True)
    out = fa.flash_attention_qkv_packed(x, 1, 1)
    out.backward(g_out)
    for n in (0, 7, 31, 32, 39):
        x1 = qkv[:, n:n + 1].clone().requires_grad_(True)
        o1 = fa.flash_attention_qkv_packed(x1, 1, 1)
        o1.backward(g_out[n:n + 1])
        assert torch.equal(o1[0], out[n]), n
        assert torch.equal(x1.grad[1, 0], x.grad[1, n]) and torch.equal(x1.grad[2, 0], x.grad[2, n]), n
        assert rel_fro(x1.grad[0, 0].float(), x.grad[0, n].float()) < 1e-3, n


@pytest.mark.parametrize("H,Hkv,D,causal,dtype,lens_q,lens_k", [
    (4, 2, 64, True, "bf16", [200, 1, 331, 128], [200, 1, 331, 128]),     # self-attention, grouped-query
    (2, 2, 128, False, "fp16", [130, 64, 257], [72, 264, 8]),             # cross-attention lengths, non-causal
    (3, 3, 40, True, "fp32", [50, 77], [50, 77]),                          # SIMT family through the same views
])
def test_varlen_packed_tokens(H, Hkv, D, causal, dtype, lens_q, lens_k):
    """flash_attention_ipu_b200.varlen: sequences packed along the token axis, one strided-view problem per
    sequence written in place -- each sequence equals fp64 attention of that sequence alone, and no token of one
    sequence influences another."""
    from flash_attention_ipu_b200.varlen import flash_attention_varlen
    torch.manual_seed(SEED + D)
    dt = DT[dtype]
    tq, tk = sum(lens_q), sum(lens_k)
    q, dout = (torch.randn(tq, H, D).to(dt) for _ in range(2))
    k, v = (torch.randn(tk, Hkv, D).to(dt) for _ in range(2))
    cq = [0] + list(np.cumsum(lens_q))
    ck = torch.tensor([0] + list(np.cumsum(lens_k)), dtype=torch.int32)   # a tensor works as well as a list
    xs = [t.cuda().requires_grad_(True) for t in (q, k, v)]
    out = flash_attention_varlen(*xs, cq, ck, causal=causal)
    out.backward(dout.cuda())
    torch.cuda.synchronize()
    tol_o, tol_g = (1e-4, 1e-4) if dtype == "fp32" else (1e-2, 1e-2)
    for s in range(len(lens_q)):
        q0, q1, k0, k1 = cq[s], cq[s + 1], int(ck[s]), int(ck[s + 1])
        bhld = lambda t, a, b: t[a:b].permute(1, 0, 2)[None].float()
        ref = _blocked_reference(bhld(q, q0, q1), bhld(k, k0, k1).repeat_interleave(H // Hkv, 1),
                                 bhld(v, k0, k1).repeat_interleave(H // Hkv, 1), bhld(dout, q0, q1), causal, 0, D ** -0.5)
        got = out.detach()[q0:q1].permute(1, 0, 2)[None].float().cpu().double()
        assert (got - ref[0]).abs().max().item() <= tol_o * 2, s
        # a one-token sequence has dq = 0 exactly (softmax of one score): absolute bound where the reference vanishes
        close = lambda a, b: rel_fro(a, b) < tol_g or (a - b).abs().max().item() < 1e-6
        assert close(xs[0].grad[q0:q1].permute(1, 0, 2)[None].float().cpu().double(), ref[2]), s
        for x, r in ((xs[1], ref[3]), (xs[2], ref[4])):
            want = r.view(1, Hkv, H // Hkv, k1 - k0, D).sum(2)            # query heads that share a K/V head
            assert close(x.grad[k0:k1].permute(1, 0, 2)[None].float().cpu().double(), want), s
    with pytest.raises(ValueError):
        flash_attention_varlen(*xs, [0, tq - 1], ck)


def test_two_streams_and_graph_replay():
    """Launch plumbing: (1) forwards + backwards queued on two streams at once (separate work-counter slots,
    programmatic dependent launch inside each stream) give the results of the serial run; (2) a CUDA graph that
    captures fa_b200_fwd + fa_b200_bwd_noalloc (no allocation, no memset inside the capture) replays to the same
    bytes."""
    native = _native_loaded()
    torch.manual_seed(SEED)
    B, H, L, D = 2, 12, 2048, 128         # 192 forward / 384 backward items per problem: more than SMs, so the atomic
    # work counters are in use; the backward scratch (25 MB) fits the no-alloc entry point's 32 MiB
    mk = lambda: tuple(torch.randn(B, H, L, D, device="cuda").bfloat16() for _ in range(4))
    probs = [mk() for _ in range(2)]

    def run(q, k, v, dout, bwd):
        prob = native.make_problem(q, k, True, D ** -0.5)
        o, lse = torch.empty_like(q), torch.empty(B, H, L, device="cuda", dtype=torch.float32)
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        native.fwd(prob, q, k, v, o, lse)
        bwd(prob, q, k, v, o, dout, lse, dq, dk, dv)
        return o, dq, dk, dv

    def with_ws(prob, q, k, v, o, dout, lse, dq, dk, dv):
        ws = torch.empty(native.bwd_workspace_bytes(prob), device="cuda", dtype=torch.uint8)
        native.bwd(prob, q, k, v, o, dout, lse, dq, dk, dv, ws)

    serial = [run(*p, with_ws) for p in probs]
    torch.cuda.synchronize()
    streams = [torch.cuda.Stream() for _ in range(2)]
    for rep in range(3):
        got = []
        for s, p in zip(streams, probs):
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                got.append(run(*p, with_ws))
        for s in streams:
            torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        for a, b in zip(got, serial):
            assert torch.equal(a[0], b[0]) and torch.equal(a[2], b[2]) and torch.equal(a[3], b[3]), rep
            assert rel_fro(a[1].float(), b[1].float()) < 1e-3, rep

    # (2) graph capture of the allocation-free pair; the warm-up call above has set the kernel attributes, and one
    # eager call of the no-alloc entry point looks up the scratch address outside the capture
    q, k, v, dout = probs[0]
    prob = native.make_problem(q, k, True, D ** -0.5)
    o, lse = torch.empty_like(q), torch.empty(B, H, L, device="cuda", dtype=torch.float32)
    dq, dk, dv = torch.zeros_like(q), torch.zeros_like(k), torch.zeros_like(v)
    native.fwd(prob, q, k, v, o, lse)
    native.bwd_noalloc(prob, q, k, v, o, dout, lse, dq, dk, dv)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            native.fwd(prob, q, k, v, o, lse)
            native.bwd_noalloc(prob, q, k, v, o, dout, lse, dq, dk, dv)
    torch.cuda.current_stream().wait_stream(side)
    for rep in range(2):
        for t in (o, dq, dk, dv):
            t.fill_(float("nan"))
        g.replay()
        torch.cuda.synchronize()
        assert torch.equal(o, serial[0][0]) and torch.equal(dk, serial[0][2]) and torch.equal(dv, serial[0][3]), rep
        assert rel_fro(dq.float(), serial[0][1].float()) < 1e-3, rep


def test_dropout_with_key_padding():
    """kv_lens and dropout together (the usual BERT-style training call: padding mask + attention dropout): the
    padded keys are outside the softmax, the surviving ones are thinned by the hash mask."""
    native = _native_loaded()
    torch.manual_seed(SEED)
    B, H, L, D, p_drop, seed = 2, 2, 384, 64, 0.2, 77
    lens = [384, 150]
    q, k, v, dout = (torch.randn(B, H, L, D).bfloat16() for _ in range(4))
    keep, drop_scale = ao.dropout_keep_mask(seed, B * H, L, L, p_drop)
    keep = keep.view(B, H, L, L)
    valid = (torch.arange(L)[None, :] < torch.tensor(lens)[:, None])            # (B, L)
    qd, kd, vd = (t.double().requires_grad_(True) for t in (q, k, v))
    s = torch.einsum("bhqd,bhkd->bhqk", qd, kd) * D ** -0.5
    s = s.masked_fill(~valid[:, None, None, :], float("-inf"))
    ref = torch.einsum("bhqk,bhkd->bhqd", torch.softmax(s, -1) * keep.double() * drop_scale, vd)
    ref.backward(dout.double())
    xs = [t.cuda() for t in (q, k, v, dout)]
    lens_t = torch.tensor(lens, dtype=torch.int32).repeat_interleave(H).cuda()
    prob = native.make_problem(xs[0], xs[1], False, D ** -0.5, kv_lens=lens_t, dropout_p=p_drop, dropout_seed=seed)
    o, lse = torch.empty_like(xs[0]), torch.empty(B, H, L, device="cuda", dtype=torch.float32)
    native.fwd(prob, xs[0], xs[1], xs[2], o, lse)
    dq, dk, dv = (torch.empty_like(t) for t in xs[:3])
    ws = torch.empty(native.bwd_workspace_bytes(prob), device="cuda", dtype=torch.uint8)
    native.bwd(prob, xs[0], xs[1], xs[2], o, xs[3], lse, dq, dk, dv, ws)
    torch.cuda.synchronize()
    assert rel_fro(o.cpu().double(), ref.detach()) < 1e-2
    for got, want in zip((dq, dk, dv), (qd.grad, kd.grad, vd.grad)):
        assert rel_fro(got.cpu().double(), want) < 1e-2
    assert torch.count_nonzero(dk[1, :, 150:]) == 0 and torch.count_nonzero(dv[1, :, 150:]) == 0
