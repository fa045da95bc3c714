"""CPU tests: the oracle is pinned against outputs of the reference's own code (tests/golden,
made by oracle/gen_golden.py) and its three restatements agree with each other."""
import numpy as np
import pytest
import torch

from conftest import GOLDEN, packed_inputs, sdpa_inputs
from oracle import attention_oracle as ao

PACKED = [("packed_g2_l256_d128", 2, 256, 128), ("packed_g3_l128_d64", 3, 128, 64), ("packed_g4_l6_d2", 4, 6, 2)]


@pytest.mark.parametrize("name,G,L,D", PACKED)
def test_python_oracle_matches_reference_outputs(name, G, L, D):
    gold = np.load(GOLDEN / f"{name}.npz")
    qkv, grad = packed_inputs(G, L, D)
    assert abs(qkv.double().sum().item() + grad.double().sum().item() - float(gold["in_checksum"])) < 1e-6
    x = qkv.clone().requires_grad_(True)
    out = ao.packed_attention_reference(x)
    (out * grad).sum().backward()
    np.testing.assert_allclose(out.detach().numpy(), gold["out"], atol=1e-6, rtol=0)
    np.testing.assert_allclose(x.grad.numpy(), gold["dqkv"], atol=2e-6, rtol=0)
    # analytic backward (vanilla_attention.cpp:68-124) against the reference's autograd result
    np.testing.assert_allclose(ao.vanilla_attention_grad(grad, qkv).numpy(), gold["dqkv"], atol=1e-5, rtol=0)


@pytest.mark.parametrize("name,G,L,D", PACKED)
def test_c_oracle_matches_reference_outputs(name, G, L, D, c_oracle):
    gold = np.load(GOLDEN / f"{name}.npz")
    qkv, grad = packed_inputs(G, L, D)
    out, lse = c_oracle.fwd(qkv.numpy())
    np.testing.assert_allclose(out, gold["out"], atol=2e-6, rtol=0)
    np.testing.assert_allclose(c_oracle.bwd(grad.numpy(), qkv.numpy()), gold["dqkv"], atol=1e-5, rtol=0)
    _, lse_py = ao.packed_attention_reference_with_lse(qkv)
    np.testing.assert_allclose(lse, lse_py.numpy(), atol=1e-5, rtol=0)


def test_sdpa_oracle_matches_reference_outputs():
    gold = np.load(GOLDEN / "sdpa_b2x3_l256_d64.npz")
    qkv, grad = sdpa_inputs((2, 3), 256, 64)
    x = qkv.clone().requires_grad_(True)
    out = ao.sdpa_reference(x[0], x[1], x[2], is_causal=True)
    (out * grad).sum().backward()
    np.testing.assert_allclose(out.detach().numpy(), gold["out"], atol=1e-6, rtol=0)
    np.testing.assert_allclose(x.grad.numpy(), gold["dqkv"], atol=1e-5, rtol=0)
    # the packed path with q pre-scaled is the same function (scaled_dot_product_attention.py:55-57)
    packed = torch.stack([qkv[0] * 64 ** -0.5, qkv[1], qkv[2]]).reshape(3, 6, 256, 64)
    np.testing.assert_allclose(ao.packed_attention_reference(packed).reshape(2, 3, 256, 64).numpy(),
                               gold["out"], atol=2e-6, rtol=0)


# chunk pairs of tests/cpp/test_flash_attention.cpp:245-259 (ck >= 2: the reference hangs at ck == 1)
@pytest.mark.parametrize("G,L,D,pairs", [
    (4, 6, 2, [(1, 1), (2, 1), (1, 2), (2, 2), (3, 2), (2, 3), (3, 3)]),
    (2, 64, 16, [(1, 1), (2, 4), (4, 2), (8, 8), (16, 2)]),
])
def test_chunked_restatement_is_chunk_invariant(G, L, D, pairs):
    torch.manual_seed(0)
    qkv = torch.randn(3, G, L, D, dtype=torch.float64)
    grad = torch.randn(G, L, D, dtype=torch.float64)
    ref = ao.packed_attention_reference(qkv)
    dref = ao.vanilla_attention_grad(grad, qkv)
    for ncq, nckv in pairs:
        out, lse = ao.chunked_attention_forward(qkv, ncq, nckv)
        assert (out - ref).abs().max() < 1e-12
        assert (ao.chunked_attention_backward(grad, qkv, ncq, nckv) - dref).abs().max() < 1e-11
        assert (lse - ao.packed_attention_reference_with_lse(qkv)[1]).abs().max() < 1e-12


def test_merge_partials_equals_full_attention():
    torch.manual_seed(1)
    q, k, v = torch.randn(3, 2, 32, 8, dtype=torch.float64)
    full = torch.softmax(q @ k.transpose(1, 2), -1) @ v
    parts = []
    for sl in (slice(0, 20), slice(20, 32)):
        s = q @ k[:, sl].transpose(1, 2)
        lse = torch.logsumexp(s, -1)
        parts.append((torch.exp(s - lse[..., None]) @ v[:, sl], lse))
    o, lse = ao.merge_partials(*parts[0], *parts[1])
    assert (o - full).abs().max() < 1e-12
    assert (lse - torch.logsumexp(q @ k.transpose(1, 2), -1)).abs().max() < 1e-12


def test_heuristic_chunks():
    # _impl/scaled_dot_product_attention.py:52-53
    assert ao.heuristic_chunks(256, 64) == (2, 2)
    assert ao.heuristic_chunks(1024, 64) == (4, 4)
    assert ao.heuristic_chunks(8192, 128) == (8, 8)
    assert ao.heuristic_chunks(512, 64) == (4, 2)
    assert ao.heuristic_chunks(64, 128) == (1, 1)


@pytest.mark.parametrize("seed,p_drop,q0,k0", [(0, 0.25, 0, 0), (0x1234_5678_9ABC_DEF1, 0.1, 128, 0),
                                               (2 ** 63 + 12345, 0.5, 0, 64), (99, 0.003, 7, 4)])
def test_dropout_mask_restatement_matches_the_library(seed, p_drop, q0, k0):
    """oracle/attention_oracle.py: dropout_keep_mask against fa_b200_dropout_keep_host, which evaluates the
    functions the kernels compile (csrc/sm100_ptx.cuh: drop_hash) on the host: pins the restatement of the
    hash, the threshold and the scale without a GPU.  (Against the REFERENCE it stays unpinned: the
    reference has no dropout.)"""
    from flash_attention_ipu_b200._impl import native
    B, H, Lq, Lk, D = 2, 3, 37, 203, 64
    prob = native.Problem(B, H, Lq, Lk, D, native.FA_B200_BF16, 1, 0.125, 0, 0, q0, k0, 0, None, p_drop,
                          seed & 0xFFFFFFFFFFFFFFFF)
    keep, scale = ao.dropout_keep_mask(seed, B * H, Lq, Lk, p_drop, q0, k0)
    for g in range(B * H):
        for r in (0, 1, Lq // 2, Lq - 1):
            got, got_scale = native.dropout_keep_host(prob, g, r)
            assert torch.equal(got.bool(), keep[g, r]), (g, r)
            assert abs(got_scale - scale) < 1e-6
    part, _ = native.dropout_keep_host(prob, 1, 5, 10, 50)
    assert torch.equal(part.bool(), keep[1, 5, 10:60])
    frac = keep.float().mean().item()
    thresh = max(1, round(p_drop * 256))
    assert abs(frac - (1 - thresh / 256)) < 0.02
    # rows, groups and seeds decorrelate
    assert not torch.equal(keep[0, 0], keep[0, 1]) and not torch.equal(keep[0], keep[1])
    other, _ = ao.dropout_keep_mask(seed + 1, B * H, Lq, Lk, p_drop, q0, k0)
    assert not torch.equal(other, keep)
