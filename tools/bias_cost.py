"""What an additive attn_mask, dropout and a key-padding mask cost: fwd+bwd of one shape through the SDPA override,
CUDA-event ms per step (median of 20), against the plain causal / non-causal call."""
import sys
import torch
import torch.nn.functional as F
sys.path.insert(0, ".")
import flash_attention_ipu.auto  # noqa: F401

B, H, T, D = 4, 32, 2048, (int(sys.argv[1]) if len(sys.argv) > 1 else 128)
q, k, v = (torch.randn(B, H, T, D, device="cuda", dtype=torch.bfloat16, requires_grad=True) for _ in range(3))
g = torch.randn(B, H, T, D, device="cuda", dtype=torch.bfloat16)
idx = torch.arange(T, device="cuda")
alibi = (-0.05 * (idx[:, None] - idx[None, :]).abs().float())[None, None].expand(1, H, T, T).bfloat16().contiguous()
causal_bool = (idx[None, :] <= idx[:, None])
pad = (idx[None, :] < torch.tensor([2048, 1500, 1024, 700], device="cuda")[:, None]).view(B, 1, 1, T)
left = torch.zeros(B, 1, 1, T, device="cuda", dtype=torch.bfloat16).masked_fill_(~pad.flip(-1), float("-inf"))  # float key mask: (B, 1, 1, T), one row for every query
flops = 4.0 * B * H * T * T * D * 3.5


def timed(fn, n=20):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[n // 2]


def step(**kw):
    def f():
        q.grad = k.grad = v.grad = None
        F.scaled_dot_product_attention(q, k, v, **kw).backward(g)
    return f


for name, kw, frac in [("non-causal", {}, 1.0), ("causal (is_causal)", {"is_causal": True}, 0.5),
                       ("float bias (1,H,T,T)", {"attn_mask": alibi}, 1.0), ("bool causal mask (T,T) as bias", {"attn_mask": causal_bool}, 1.0),
                       ("key padding (kv_lens)", {"attn_mask": pad}, 1.0), ("left padding (key mask as bias)", {"attn_mask": left}, 1.0),
                       ("dropout 0.1", {"dropout_p": 0.1}, 1.0),
                       ("dropout 0.1 causal", {"dropout_p": 0.1, "is_causal": True}, 0.5)]:
    ms = timed(step(**kw))
    print(f"{name:34s} {ms:7.3f} ms  {flops * frac / ms / 1e9:7.1f} TFLOP/s (of the dense count)")

# forward alone (no autograd), with and without the float bias
with torch.no_grad():
    for name, kw in [("forward only, no mask", {}), ("forward only, float bias", {"attn_mask": alibi})]:
        ms = timed(lambda: F.scaled_dot_product_attention(q, k, v, **kw))
        print(f"{name:34s} {ms:7.3f} ms")
