"""Per-chunk timeline (CUDA events) of one HostStreamedAttention step in steady state."""
import sys, torch
sys.path.insert(0, ".")
import flash_attention_ipu_b200.streaming as st
from flash_attention_ipu_b200.streaming import HostStreamedAttention
import os
G, T, D, gc = int(os.environ.get("E2E_G", "128")), 8192, 128, 8
h_qkv = torch.randn(3, G, T, D).bfloat16().pin_memory()
h_do = torch.randn(G, T, D).bfloat16().pin_memory()
h_out = torch.empty(G, T, D, dtype=torch.bfloat16).pin_memory()
h_dqkv = torch.empty(3, G, T, D, dtype=torch.bfloat16).pin_memory()
marks = []
def ev(stream, tag):
    e = torch.cuda.Event(enable_timing=True); e.record(stream); marks.append((tag, e))
class Traced(HostStreamedAttention):
    def run(self, h_qkv, h_dout, h_out, h_dqkv, sync=False):
        cur = torch.cuda.current_stream(self.dev)
        gc = self.gc
        for c in range(self.n + 1):
            s = c % 2
            if c < self.n:
                with torch.cuda.stream(self.s_in):
                    self.s_in.wait_event(self.ev_free[s])
                    ev(self.s_in, f"up{c}+")
                    sl = slice(c * gc, (c + 1) * gc)
                    for i in range(3):
                        self.d_qkv[s][i].copy_(h_qkv[i, sl], non_blocking=True)
                    self.d_do[s].copy_(h_dout[sl], non_blocking=True)
                    ev(self.s_in, f"up{c}-")
                    self.ev_in[s].record(self.s_in)
            if c >= 1:
                p = (c - 1) % 2
                cur.wait_event(self.ev_in[p])
                ev(cur, f"cmp{c-1}+")
                x = self.d_qkv[p].detach().requires_grad_(True)
                out = st.flash_attention_qkv_packed(x, self.ncq, self.nckv, causal=self.causal, softmax_scale=self.scale)
                out.backward(self.d_do[p])
                ev(cur, f"cmp{c-1}-")
                self.ev_free[p].record(cur); self.ev_done[p].record(cur)
                self.res[p] = (out.detach(), x.grad)
                with torch.cuda.stream(self.s_out):
                    self.s_out.wait_event(self.ev_done[p])
                    ev(self.s_out, f"dn{c-1}+")
                    sl = slice((c - 1) * gc, c * gc)
                    h_out[sl].copy_(self.res[p][0], non_blocking=True)
                    for i in range(3):
                        h_dqkv[i, sl].copy_(self.res[p][1][i], non_blocking=True)
                    ev(self.s_out, f"dn{c-1}-")
                    self.res[p][0].record_stream(self.s_out); self.res[p][1].record_stream(self.s_out)
        cur.wait_stream(self.s_out); cur.wait_stream(self.s_in)
pipe = Traced(G, T, D, torch.bfloat16, torch.device("cuda"), groups_per_chunk=gc, num_chunks_q=8, num_chunks_kv=8)
for _ in range(2): pipe.run(h_qkv, h_do, h_out, h_dqkv)
torch.cuda.synchronize(); marks.clear()
pipe.run(h_qkv, h_do, h_out, h_dqkv); torch.cuda.synchronize()
t0 = marks[0][1]
rows = {}
for tag, e in marks:
    rows.setdefault(tag[:-1], {})[tag[-1]] = t0.elapsed_time(e)
tag = os.environ.get("E2E_TAG", "")
for k, v in rows.items():
    print(f"{tag}{k:6s} {v['+']:8.2f} -> {v['-']:8.2f}  ({v['-'] - v['+']:.2f} ms)")
import time
t = time.perf_counter()
for _ in range(5): pipe.run(h_qkv, h_do, h_out, h_dqkv); torch.cuda.synchronize()
print(f"{tag}step wall {(time.perf_counter() - t) / 5 * 1e3:.2f} ms for G={G}")
