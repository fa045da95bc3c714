"""Timing of the deterministic-dQ backward for a few shapes (debug aid)."""
import sys, time, torch
sys.path.insert(0, ".")
import flash_attention_ipu as fa, flash_attention_ipu_b200 as pkg
for G, L in ((1, 8192), (2, 8192), (16, 8192), (128, 8192)):
    qkv = torch.randn(3, G, L, 128, device="cuda", dtype=torch.bfloat16).requires_grad_(True)
    g = torch.randn(G, L, 128, device="cuda", dtype=torch.bfloat16)
    for det in (False, True):
        pkg.set_deterministic(det)
        for _ in range(2):
            qkv.grad = None; fa.flash_attention_qkv_packed(qkv, 8, 8).backward(g)
        torch.cuda.synchronize(); t = time.perf_counter()
        for _ in range(5):
            qkv.grad = None; fa.flash_attention_qkv_packed(qkv, 8, 8).backward(g)
        torch.cuda.synchronize()
        print(f"G={G} L={L} {'deterministic' if det else 'default':13s} {(time.perf_counter() - t) / 5 * 1e3:8.3f} ms", flush=True)
pkg.set_deterministic(False)
