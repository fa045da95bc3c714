"""Why is the e2e leg slower under torchrun at N > 1?  Times HostStreamedAttention.run per step in three
settings inside one torchrun launch: before init_process_group, after it, and with sync=True."""
import os, sys, time, torch
sys.path.insert(0, ".")
rank = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(rank)
dev = torch.device("cuda", rank)
from flash_attention_ipu_b200.streaming import HostStreamedAttention
G, T, D = int(os.environ.get("E2E_G", "64")), 8192, 128
h_qkv = torch.randn(3, G, T, D).bfloat16().pin_memory()
h_do = torch.randn(G, T, D).bfloat16().pin_memory()
h_out = torch.empty(G, T, D, dtype=torch.bfloat16).pin_memory()
h_dqkv = torch.empty(3, G, T, D, dtype=torch.bfloat16).pin_memory()
pipe = HostStreamedAttention(G, T, D, torch.bfloat16, dev, groups_per_chunk=8, num_chunks_q=8, num_chunks_kv=8)
def timed(tag, sync):
    for _ in range(2): pipe.run(h_qkv, h_do, h_out, h_dqkv, sync=sync)
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(5): pipe.run(h_qkv, h_do, h_out, h_dqkv, sync=sync)
    torch.cuda.synchronize()
    print(f"rank {rank} {tag}: {(time.perf_counter() - t) / 5 * 1e3:.2f} ms/step (OMP_NUM_THREADS={os.environ.get('OMP_NUM_THREADS')}, torch threads {torch.get_num_threads()})", flush=True)
timed("no process group, sync=False", False)
timed("no process group, sync=True", True)
if "RANK" in os.environ:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)
    dist.barrier(); torch.cuda.synchronize()
    timed("after init_process_group(nccl), sync=True", True)
    dist.barrier()
    dist.destroy_process_group()
