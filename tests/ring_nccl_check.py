#!/usr/bin/env python
"""Multi-GPU check of the NCCL ring (run by hand / by scripts/gpu_ring.sh, not collected by pytest):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29544 tests/ring_nccl_check.py [T]

Every rank builds the same full (1, 4, T, 128) bf16 problem from one seed, runs
``ring_attention`` on its zig-zag shard over real NCCL P2P, and compares output and gradients with
the un-sharded CUDA kernel run locally on the full sequence.  Prints one line per rank."""
import os
import pathlib
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))


def main():
    T = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)
    from flash_attention_ipu_b200 import ring
    import flash_attention_ipu.auto  # noqa: F401
    torch.manual_seed(1123581321)
    q, k, v, dout = (torch.randn(1, 4, T, 128).bfloat16().to(dev) for _ in range(4))
    ql, kl, vl = (ring.zigzag_shard(t, rank, world).requires_grad_(True) for t in (q, k, v))
    out = ring.ring_attention(ql, kl, vl)
    out.backward(ring.zigzag_shard(dout, rank, world))
    qf, kf, vf = (t.clone().requires_grad_(True) for t in (q, k, v))
    ref = torch.nn.functional.scaled_dot_product_attention(qf, kf, vf, is_causal=True)
    ref.backward(dout)
    torch.cuda.synchronize()
    sh = lambda t: ring.zigzag_shard(t, rank, world).float()
    rel = lambda a, b: ((a.float() - b).norm() / b.norm()).item()
    errs = dict(out_maxabs=(out.float() - sh(ref.detach())).abs().max().item(), dq=rel(ql.grad, sh(qf.grad)),
                dk=rel(kl.grad, sh(kf.grad)), dv=rel(vl.grad, sh(vf.grad)))
    ok = errs["out_maxabs"] <= 1e-2 and max(errs["dq"], errs["dk"], errs["dv"]) <= 1e-2
    print(f"rank {rank}/{world} T={T} {'OK' if ok else 'FAIL'} " + " ".join(f"{k}={v:.2e}" for k, v in errs.items()),
          flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
