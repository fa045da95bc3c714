"""CPU test (gloo, world_size 2) of the multi-GPU host logic: (batch, head) groups are split
across ranks with no data-path collective, every group is processed exactly once, and the
max-over-ranks timing reduction that bench.py uses works.  The per-group computation here is the
CPU oracle standing in for the CUDA op (host-logic test only; the CUDA path is tested with -m gpu)."""
import os
import sys
import pathlib

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def _load_sharding():
    """The product's sharding module, loaded by path: importing the package itself needs the native
    library, which is not what this CPU test is about."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("fa_sharding", ROOT / "flash_attention_ipu_b200" / "sharding.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


shard_groups = _load_sharding().shard_groups


def _worker(rank, world, port, G, L, D, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import attention_oracle as ao
    torch.manual_seed(0)
    qkv = torch.randn(3, G, L, D)
    lo, hi = shard_groups(G, world, rank)
    assert _load_sharding().local_groups(qkv, world, rank).shape[1] == hi - lo
    out_local = ao.packed_attention_reference(_load_sharding().local_groups(qkv, world, rank))
    # no collective on the data path: each rank writes its own shard; only timing is reduced
    torch.save((lo, hi, out_local), f"{tmp}/shard{rank}.pt")
    t = torch.tensor([float(rank + 1)])
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    assert t.item() == world
    dist.barrier()
    dist.destroy_process_group()


def test_group_sharding_two_ranks(tmp_path):
    G, L, D, world = 5, 32, 8, 2
    port = 29500 + os.getpid() % 1000
    mp.spawn(_worker, args=(world, port, G, L, D, str(tmp_path)), nprocs=world, join=True)
    from oracle import attention_oracle as ao
    torch.manual_seed(0)
    qkv = torch.randn(3, G, L, D)
    full = ao.packed_attention_reference(qkv)
    covered = torch.zeros(G, dtype=torch.int)
    for r in range(world):
        lo, hi, out = torch.load(f"{tmp_path}/shard{r}.pt")
        covered[lo:hi] += 1
        assert torch.equal(out, full[lo:hi])
    assert (covered == 1).all()


def test_shard_groups_balanced():
    for G in (1, 7, 128, 129):
        for world in (1, 2, 4, 8):
            spans = [shard_groups(G, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == G
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
