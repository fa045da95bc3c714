"""Host-side logic of the sequence-sharded ring (flash_attention_ipu_b200/ring.py) on CPU.

The per-block kernels are replaced by plain fp64 torch math with the same three methods
(``OracleBlocks`` -- test infrastructure, never used by the product), so that what is tested is
the zig-zag layout, the one-block-per-hop schedule, the fp32 LSE-merge recurrence, the travelling
dK/dV accumulators and the real P2P exchange (gloo, world_size 2 and 3).  The CUDA blocks are
tested with ``-m gpu`` in test_gpu_parity.py (emulated ring vs the single-GPU kernel / oracle).
"""
import os
import pathlib
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def _ring_module():
    # ring.py lives in the product package, whose __init__ needs the built library (present after
    # `make`); the host logic itself is import-light
    from flash_attention_ipu_b200 import ring
    return ring


class OracleBlocks:
    """fp64 restatement of one block (cpp/flash_attention_qkv_packed.cpp:254-300 forward,
    :457-472 backward with the GLOBAL lse), standing in for the CUDA kernels."""

    @staticmethod
    def _scores(q, k, causal, scale):
        s = torch.einsum("bhqd,bhkd->bhqk", q.double(), k.double()) * scale
        if causal:
            Lq, Lk = s.shape[-2:]
            s = s.masked_fill(torch.ones(Lq, Lk, dtype=torch.bool).triu(1), float("-inf"))
        return s

    def fwd_block(self, q, k, v, causal, scale):
        s = self._scores(q, k, causal, scale)
        lse = torch.logsumexp(s, dim=-1)
        o = torch.einsum("bhqk,bhkd->bhqd", torch.exp(s - lse[..., None]), v.double())
        return o.to(q.dtype), lse.float()

    def bwd_block(self, dout, q, k, v, out, lse, causal, scale):
        s = self._scores(q, k, causal, scale)
        p = torch.exp(s - lse.double()[..., None])
        delta = (out.double() * dout.double()).sum(-1, keepdim=True)
        dv = torch.einsum("bhqk,bhqd->bhkd", p, dout.double())
        dp = torch.einsum("bhqd,bhkd->bhqk", dout.double(), v.double())
        ds = p * (dp - delta) * scale
        dq = torch.einsum("bhqk,bhkd->bhqd", ds, k.double())
        dk = torch.einsum("bhqk,bhqd->bhkd", ds, q.double())
        return dq.to(q.dtype), dk.to(q.dtype), dv.to(q.dtype)

    def merge_acc(self, o_acc, lse_acc, o_b, lse_b, row0):
        n = o_b.shape[2]
        la, oa = lse_acc[:, :, row0:row0 + n], o_acc[:, :, row0:row0 + n]
        lse = torch.logaddexp(la, lse_b)
        wa = torch.where(torch.isinf(la) & (la < 0), torch.zeros_like(la), torch.exp(la - lse))
        oa.copy_(oa * wa[..., None] + o_b.float() * torch.exp(lse_b - lse)[..., None])
        la.copy_(lse)


def _full_reference(q, k, v, dout):
    q, k, v = (t.double().requires_grad_(True) for t in (q, k, v))
    o = torch.nn.functional.scaled_dot_product_attention(q, k, v, is_causal=True)
    o.backward(dout.double())
    return o.detach(), q.grad, k.grad, v.grad


def test_zigzag_roundtrip():
    ring = _ring_module()
    x = torch.arange(2 * 3 * 24 * 4, dtype=torch.float32).reshape(2, 3, 24, 4)
    for world in (1, 2, 3, 4):
        shards = [ring.zigzag_shard(x, r, world) for r in range(world)]
        assert all(s.shape[2] == 24 // world for s in shards)
        assert torch.equal(ring.zigzag_unshard(shards), x)
    with pytest.raises(ValueError):
        ring.zigzag_shard(x, 0, 5)


def test_hop_plan_covers_causal_triangle_once():
    """Every (query chunk, key chunk) pair at or below the diagonal is computed by exactly one
    hop of exactly one rank, nothing above the diagonal is, and every hop costs 2 chunk-blocks."""
    ring = _ring_module()
    for world in (1, 2, 3, 4, 8):
        n = 2 * world
        seen = torch.zeros(n, n, dtype=torch.int)
        for r in range(world):
            own = {"lo": [r], "hi": [n - 1 - r], "all": [r, n - 1 - r]}
            for s in range(world):
                src = (r - s) % world
                theirs = {"lo": [src], "hi": [n - 1 - src], "all": [src, n - 1 - src]}
                qr, kr, causal = ring.hop_plan(r, src)
                cost = 0.0
                for a in own[qr]:
                    for b in theirs[kr]:
                        if causal and b > a:
                            continue  # masked quarter of the local block
                        seen[a, b] += 1
                        cost += 0.5 if a == b else 1.0
                assert cost == 2.0
        assert torch.equal(seen, torch.ones(n, n, dtype=torch.int).tril())


@pytest.mark.parametrize("world", [1, 2, 4])
def test_emulated_ring_matches_full_attention(world):
    ring = _ring_module()
    torch.manual_seed(1123581321)
    B, H, T, D = 1, 2, 16 * world, 8
    q, k, v, dout = (torch.randn(B, H, T, D) for _ in range(4))
    sh = lambda t: [ring.zigzag_shard(t, r, world) for r in range(world)]
    outs, grads = ring.ring_attention_emulated(sh(q), sh(k), sh(v), sh(dout), blocks=OracleBlocks())
    o_ref, dq_ref, dk_ref, dv_ref = _full_reference(q, k, v, dout)
    assert torch.allclose(ring.zigzag_unshard(outs).double(), o_ref, atol=2e-6)
    for i, ref in enumerate((dq_ref, dk_ref, dv_ref)):
        got = ring.zigzag_unshard([g[i] for g in grads]).double()
        assert torch.allclose(got, ref, atol=5e-6), (i, (got - ref).abs().max())


def _worker(rank, world, port, T, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ring = _ring_module()
    torch.manual_seed(7)
    q, k, v, dout = (torch.randn(1, 2, T, 8) for _ in range(4))
    ql, kl, vl = (ring.zigzag_shard(t, rank, world).requires_grad_(True) for t in (q, k, v))
    out = ring.ring_attention(ql, kl, vl, blocks=OracleBlocks())
    out.backward(ring.zigzag_shard(dout, rank, world))
    torch.save((out.detach(), ql.grad, kl.grad, vl.grad), f"{tmp}/rank{rank}.pt")
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_ring_p2p_gloo(world, tmp_path):
    ring = _ring_module()
    T = 12 * world
    port = 29700 + (os.getpid() + world) % 1000
    mp.spawn(_worker, args=(world, port, T, str(tmp_path)), nprocs=world, join=True)
    torch.manual_seed(7)
    q, k, v, dout = (torch.randn(1, 2, T, 8) for _ in range(4))
    refs = _full_reference(q, k, v, dout)
    parts = [torch.load(f"{tmp_path}/rank{r}.pt") for r in range(world)]
    for i, ref in enumerate(refs):
        got = ring.zigzag_unshard([p[i] for p in parts]).double()
        assert torch.allclose(got, ref, atol=5e-6), (i, (got - ref).abs().max())


def test_ring_refuses_cpu_tensors_without_injected_blocks():
    ring = _ring_module()
    x = torch.randn(1, 1, 8, 8)
    with pytest.raises(RuntimeError):
        ring.ring_attention(x, x, x)
