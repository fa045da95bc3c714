"""One fwd+bwd step of the GPT-2-small and BERT-large attention shapes through the SDPA override, for an
ncu launch list (`ncu --metrics gpu__time_duration.sum`): which kernel the small shapes spend their time in."""
import sys
import torch
import torch.nn.functional as F
sys.path.insert(0, ".")
import flash_attention_ipu.auto  # noqa: F401

for name, (B, H, T, D, causal) in {"gpt2": (8, 12, 1024, 64, True), "bert": (32, 16, 512, 64, False)}.items():
    q, k, v = (torch.randn(B, H, T, D, device="cuda", dtype=torch.bfloat16, requires_grad=True) for _ in range(3))
    g = torch.randn(B, H, T, D, device="cuda", dtype=torch.bfloat16)
    for _ in range(3):
        torch.cuda.nvtx.range_push(name)
        out = F.scaled_dot_product_attention(q, k, v, is_causal=causal)
        out.backward(g)
        torch.cuda.nvtx.range_pop()
    torch.cuda.synchronize()
print("ok")
