"""Where the e2e step time goes: HostStreamedAttention at several chunk sizes, timed over 5
back-to-back steps (as bench.py does), against the duplex PCIe floor of tools/pcie_probe.py."""
import sys, torch
sys.path.insert(0, ".")
from flash_attention_ipu_b200.streaming import HostStreamedAttention
G, T, D = 128, 8192, 128
h_qkv = torch.randn(3, G, T, D).bfloat16().pin_memory()
h_do = torch.randn(G, T, D).bfloat16().pin_memory()
h_out = torch.empty(G, T, D, dtype=torch.bfloat16).pin_memory()
h_dqkv = torch.empty(3, G, T, D, dtype=torch.bfloat16).pin_memory()
for gc in (64, 32, 16, 8):
    pipe = HostStreamedAttention(G, T, D, torch.bfloat16, torch.device("cuda"), groups_per_chunk=gc, num_chunks_q=8, num_chunks_kv=8)
    pipe.run(h_qkv, h_do, h_out, h_dqkv); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): pipe.run(h_qkv, h_do, h_out, h_dqkv)
    e1.record(); torch.cuda.synchronize()
    print(f"groups_per_chunk={gc}: {e0.elapsed_time(e1) / 5:.2f} ms/step")
    del pipe

# same pipeline with the attention replaced by a trivial differentiable op: the copy-only cost
import flash_attention_ipu_b200.streaming as st
st.flash_attention_qkv_packed = lambda x, *a, **k: x.sum(0)
for gc in (64, 32, 16, 8):
    pipe = HostStreamedAttention(G, T, D, torch.bfloat16, torch.device("cuda"), groups_per_chunk=gc)
    pipe.run(h_qkv, h_do, h_out, h_dqkv); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): pipe.run(h_qkv, h_do, h_out, h_dqkv)
    e1.record(); torch.cuda.synchronize()
    print(f"copy-only pipeline, groups_per_chunk={gc}: {e0.elapsed_time(e1) / 5:.2f} ms/step")
