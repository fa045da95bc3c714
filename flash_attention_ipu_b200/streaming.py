"""Host-resident attention: forward + backward of ``flash_attention_qkv_packed`` for inputs that
live in (pinned) HOST memory, streamed through the GPU in chunks of (batch, head) groups.

The groups of the packed ``(3, G, L, D)`` tensor are independent problems (the reference flattens
batch and heads into G, ``_impl/scaled_dot_product_attention.py:57``), so a step whose operands
start and end on the host does not have to be "copy everything in, compute, copy everything out":
chunk c+1 is uploaded and chunk c-1 is downloaded while chunk c computes.  PCIe is full duplex, so
the step costs about max(H2D, D2H) instead of H2D + compute + D2H.  Three CUDA streams, two
device buffers per direction, events for the hand-offs; the compute is the ordinary public op.

This is the call ``bench.py`` times for its ``e2e`` number.
"""
from typing import Optional

import torch

from ._impl.flash_attention_qkv_packed import flash_attention_qkv_packed

__all__ = ["HostStreamedAttention"]


class HostStreamedAttention:
    """Reusable pipeline object (device staging buffers and streams are allocated once).

    ``run(h_qkv, h_dout, h_out, h_dqkv)``: all four are pinned host tensors, ``h_qkv``/``h_dqkv``
    of shape (3, G, L, D) and ``h_dout``/``h_out`` of shape (G, L, D).  With ``sync=True`` (default)
    the call returns once the last download has landed: ``h_out`` holds the attention output and
    ``h_dqkv`` the gradients.  With ``sync=False`` the work is only ordered on the current CUDA stream
    (so that the next step's uploads overlap this step's tail) and the caller must synchronise that
    stream before reading the host buffers.
    """

    def __init__(self, G: int, L: int, D: int, dtype=torch.bfloat16, device: Optional[torch.device] = None,
                 groups_per_chunk: int = 16, num_chunks_q: int = 1, num_chunks_kv: int = 1,
                 causal: bool = True, softmax_scale: float = 1.0, ramp: bool = False):
        if G % groups_per_chunk:
            raise ValueError(f"G={G} is not a multiple of groups_per_chunk={groups_per_chunk}")
        self.dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        self.G, self.gc = G, groups_per_chunk
        # chunk schedule: (first group, groups).  The pipeline's fill is the upload of the first chunk and its
        # drain the compute + download of the last one; with ramp=True both ends use smaller chunks
        # (1, 1, 2, ... groups_per_chunk ..., 2, 1, 1) so that less of the step runs un-overlapped.
        sizes = [groups_per_chunk] * (G // groups_per_chunk)
        if ramp and groups_per_chunk >= 4 and len(sizes) >= 3:
            half = groups_per_chunk // 2
            head = [1, 1] + ([2] * ((half - 2) // 2) if half > 2 else []) + [groups_per_chunk - half]
            head = head if sum(head) == groups_per_chunk else [1] * groups_per_chunk
            sizes = head + sizes[1:-1] + head[::-1]
        self.chunks, g0 = [], 0
        for n_g in sizes:
            self.chunks.append((g0, n_g))
            g0 += n_g
        assert g0 == G
        self.n = len(self.chunks)
        self.ncq, self.nckv, self.causal, self.scale = num_chunks_q, num_chunks_kv, causal, softmax_scale
        mk = lambda *shape: torch.empty(shape, dtype=dtype, device=self.dev)
        # double-buffered staging: inputs (qkv, dout) and results (out, dqkv)
        self.d_qkv = [mk(3, self.gc, L, D) for _ in range(2)]
        self.d_do = [mk(self.gc, L, D) for _ in range(2)]
        self.s_in, self.s_out = torch.cuda.Stream(self.dev), torch.cuda.Stream(self.dev)
        self.ev_in = [torch.cuda.Event() for _ in range(2)]        # chunk uploaded
        self.ev_free = [torch.cuda.Event() for _ in range(2)]      # staging inputs consumed by compute
        self.ev_done = [torch.cuda.Event() for _ in range(2)]      # results of the chunk ready
        self.res = [None, None]

    def run(self, h_qkv, h_dout, h_out, h_dqkv, sync: bool = True) -> None:
        for t in (h_qkv, h_dout, h_out, h_dqkv):
            if t.is_cuda or not t.is_pinned():
                raise ValueError("HostStreamedAttention.run wants pinned host tensors")
        cur = torch.cuda.current_stream(self.dev)
        for c in range(self.n + 1):
            s = c % 2
            if c < self.n:  # ---- upload chunk c
                with torch.cuda.stream(self.s_in):
                    # the staging buffer's previous user (two chunks back, possibly the tail of the
                    # previous run) must have consumed it; a never-recorded event is a no-op
                    self.s_in.wait_event(self.ev_free[s])
                    g0, ng = self.chunks[c]
                    sl = slice(g0, g0 + ng)
                    for i in range(3):  # each [i, sl] is one contiguous block on the host
                        self.d_qkv[s][i, :ng].copy_(h_qkv[i, sl], non_blocking=True)
                    self.d_do[s][:ng].copy_(h_dout[sl], non_blocking=True)
                    self.ev_in[s].record(self.s_in)
            if c >= 1:  # ---- compute chunk c-1 (public op), then hand its results to the download stream
                p = (c - 1) % 2
                cur.wait_event(self.ev_in[p])
                g0, ng = self.chunks[c - 1]
                x = self.d_qkv[p][:, :ng].detach().requires_grad_(True)
                out = flash_attention_qkv_packed(x, self.ncq, self.nckv, causal=self.causal, softmax_scale=self.scale)
                out.backward(self.d_do[p][:ng])
                self.ev_free[p].record(cur)
                self.ev_done[p].record(cur)
                self.res[p] = (out.detach(), x.grad)
                with torch.cuda.stream(self.s_out):
                    self.s_out.wait_event(self.ev_done[p])
                    sl = slice(g0, g0 + ng)
                    h_out[sl].copy_(self.res[p][0], non_blocking=True)
                    for i in range(3):
                        h_dqkv[i, sl].copy_(self.res[p][1][i], non_blocking=True)
                    # the result tensors are fresh allocations per chunk: tell the caching allocator
                    # that the download stream still reads them
                    self.res[p][0].record_stream(self.s_out)
                    self.res[p][1].record_stream(self.s_out)
        cur.wait_stream(self.s_out)
        cur.wait_stream(self.s_in)
        if sync:
            done = torch.cuda.Event()
            done.record(cur)
            done.synchronize()
