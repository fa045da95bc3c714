#!/usr/bin/env python
"""bench.py -- causal attention fwd+bwd throughput (BASELINE.json's metric) on B200.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU math (oracle port)

One "step" = one forward + one backward of `flash_attention_qkv_packed` over the workload.
Workload at every N: BASELINE.json configs[2] (Llama-style B=4, T=8192, 32 heads, d=128, bf16,
causal) PER GPU -- the (batch, head) groups are independent, so ranks get their own groups with no
data-path collective ("scaling": "weak"; global batch = 4 N).  Inputs (805 MB qkv per rank) exceed
the 126 MB L2, so no explicit flush is needed between iterations.

Prints ONE JSON line (rank 0).  `value` times device-resident inputs with CUDA events, max over
ranks; `e2e` goes through the same public API from pinned HOST buffers with the H2D/D2H copies
inside the timed region; `roofline` is the dominant kernel against the measured cuBLAS bf16
peak in MEASURED_PEAKS.json; `cpu_baseline` times the oracle port on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
import pathlib

import torch

ROOT = pathlib.Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

WORKLOAD = dict(name="llama_attn_B4_T8192_H32_d128_bf16_causal", B=4, H=32, T=8192, D=128)
METRIC = "causal_attn_fwd_bwd_tflops"
SEED = 1123581321


def attention_flops(G, L, D, causal=True, fwd=True, bwd=True):
    f = 4.0 * G * L * L * D * (0.5 if causal else 1.0)
    return f * ((1.0 if fwd else 0.0) + (2.5 if bwd else 0.0))


def measured_peaks():
    try:
        pk = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
        return pk, "measured"
    except Exception:
        return {"bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "hbm_gbs": 6650.0}, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.proc = index, [], None

    def run(self):
        # one long-lived nvidia-smi in loop mode (a fresh process per sample is too slow for
        # timed regions of a few hundred ms)
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "20"], stdout=subprocess.PIPE, text=True)
            for line in self.proc.stdout:
                self.samples.append([x.strip() for x in line.strip().split(",")])
        except Exception:
            pass

    def mark(self):
        """Index of the next sample: call at the start and the end of the timed region."""
        return len(self.samples)

    def summary(self, lo=0, hi=None):
        """Median SM clock / throttle reasons over samples[lo:hi] (the timed region; the sampler is
        started earlier so that nvidia-smi is already running when the region begins)."""
        if self.proc is not None:
            self.proc.terminate()
        self.join(timeout=6)
        window = self.samples[lo:hi] if len(self.samples[lo:hi]) >= 3 else self.samples
        self.samples = window
        mhz = sorted(int(s[0]) for s in self.samples if len(s) >= 7 and s[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for s in self.samples if len(s) >= 7 for n, v in zip(names, s[3:7]) if v == "Active"})
        mx = max((int(s[1]) for s in self.samples if len(s) >= 7 and s[1].isdigit()), default=0)
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else 0, "sm_max_mhz": mx, "reasons": reasons,
                "samples": len(self.samples)}


def cpu_reference_step(G_sample, T, D, threads):
    """The reference's CPU math (oracle port of _impl/flash_attention_qkv_packed.py:48-53 + autograd)
    on a bounded sample of the workload: `G_sample` (batch, head) groups at full T, fp32."""
    from oracle import attention_oracle as ao
    torch.set_num_threads(threads)
    torch.manual_seed(SEED)
    qkv = torch.randn(3, G_sample, T, D)
    qkv[0] *= D ** -0.5
    grad = torch.randn(G_sample, T, D)
    x = qkv.requires_grad_(True)
    t0 = time.perf_counter()
    out = ao.packed_attention_reference(x)
    out.backward(grad)
    dt = time.perf_counter() - t0
    return dt, attention_flops(G_sample, T, D)


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the host cores.
    The Poplar C++ cannot be built here (no Poplar SDK) so this is the oracle port (kind "port")."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    G_sample = 2
    w = WORKLOAD
    for _ in range(max(args.warmup, 1) if args.warmup else 0):
        cpu_reference_step(1, w["T"], w["D"], threads)
    t_total, f_total = 0.0, 0.0
    for _ in range(args.steps):
        dt, fl = cpu_reference_step(G_sample, w["T"], w["D"], threads)
        t_total += dt
        f_total += fl
    tflops = f_total / t_total / 1e12
    sample = f"{G_sample} of {w['B'] * w['H']} (batch,head) groups per step at full T={w['T']}, d={w['D']}, fp32, torch CPU"
    line = {
        "impl": "reference", "metric": METRIC, "value": tflops, "unit": "TFLOP/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": w["name"], "sample": sample},
        "cpu_baseline": {"value": tflops, "unit": "TFLOP/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": tflops, "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: this framework has no CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    import flash_attention_ipu as fa
    from flash_attention_ipu_b200._impl import native, ops
    from flash_attention_ipu_b200._impl.scaled_dot_product_attention import _chunk_heuristic

    w = WORKLOAD
    G, T, D = w["B"] * w["H"], w["T"], w["D"]
    ncq, nckv = _chunk_heuristic(T, D)
    torch.manual_seed(SEED + rank)
    # pinned host copies of one step's inputs (for the e2e leg) and device-resident inputs
    h_qkv = torch.randn(3, G, T, D).mul_(torch.tensor([D ** -0.5, 1, 1]).view(3, 1, 1, 1)).bfloat16().pin_memory()
    h_do = torch.randn(G, T, D).bfloat16().pin_memory()
    h_out = torch.empty(G, T, D, dtype=torch.bfloat16).pin_memory()
    h_dqkv = torch.empty(3, G, T, D, dtype=torch.bfloat16).pin_memory()
    qkv = h_qkv.to(dev).requires_grad_(True)
    dout = h_do.to(dev)
    flops = attention_flops(G, T, D)

    def step():
        qkv.grad = None
        out = fa.flash_attention_qkv_packed(qkv, ncq, nckv)
        out.backward(dout)
        return out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()

    # ---- timed region: device-resident inputs, CUDA events on the launch stream
    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.3)
    launches0 = native.launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    barrier()
    m0 = sampler.mark()
    ev[0].record()
    for _ in range(args.steps):
        step()
    ev[1].record()
    barrier()
    m1 = sampler.mark()
    ms_total = ev[0].elapsed_time(ev[1])
    launches = native.launch_count() - launches0
    clocks = sampler.summary(m0, m1)
    if world > 1:
        t = torch.tensor([ms_total], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = t.item()
    ms_step = ms_total / args.steps
    value = world * flops / (ms_step * 1e-3) / 1e12

    # ---- per-kernel timing for the roofline (forward op and backward op separately)
    def time_op(fn, n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    with torch.no_grad():
        qd = qkv.detach()
        out, lse = ops.qkv_packed_fwd(qd, 1.0, True, T // ncq, T // nckv)
        n = max(3, min(args.steps, 20))
        ms_fwd = time_op(lambda: ops.qkv_packed_fwd(qd, 1.0, True, T // ncq, T // nckv), n)
        ms_bwd = time_op(lambda: ops.qkv_packed_bwd(dout, qd, out, lse, 1.0, True, T // ncq, T // nckv), n)
    peaks, peak_kind = measured_peaks()
    peak = peaks["bf16_tflops_sustained"]
    fwd_tf = attention_flops(G, T, D, bwd=False) / (ms_fwd * 1e-3) / 1e12
    bwd_tf = attention_flops(G, T, D, fwd=False) / (ms_bwd * 1e-3) / 1e12
    dominant = "bwd" if ms_bwd >= ms_fwd else "fwd"
    ach = bwd_tf if dominant == "bwd" else fwd_tf
    traffic = None  # dram bytes read + written per launch of the dominant kernel, from the committed ncu capture
    try:
        tj = json.loads((ROOT / "profiles" / "ncu_traffic.json").read_text())[f"fa_{dominant}_sm100"]
        traffic = tj["dram_bytes_read"] + tj["dram_bytes_write"]
    except Exception:
        pass
    roofline = {"bound": "tensor", "kernel": f"fa_{dominant}_sm100", "achieved": ach, "peak": peak,
                "unit": "TFLOP/s", "frac": ach / peak, "traffic": traffic,
                "peak_kind": f"{peak_kind} cuBLAS bf16 sustained (burst {peaks['bf16_tflops']})",
                "fwd_tflops": fwd_tf, "bwd_tflops": bwd_tf, "ms_fwd": ms_fwd, "ms_bwd": ms_bwd,
                "frac_of_nominal_2250": value / world / 2250.0}

    # ---- e2e: same public API, inputs start in pinned host memory, results end there.  The step
    # is pipelined over chunks of (batch, head) groups (flash_attention_ipu_b200/streaming.py):
    # upload of chunk c+1 and download of chunk c-1 overlap the attention of chunk c.
    from flash_attention_ipu_b200.streaming import HostStreamedAttention
    pipe = HostStreamedAttention(G, T, D, torch.bfloat16, dev, groups_per_chunk=8, num_chunks_q=ncq, num_chunks_kv=nckv)

    def e2e_step():
        pipe.run(h_qkv, h_do, h_out, h_dqkv)

    e2e_step()
    barrier()
    n_e2e = max(2, min(args.steps, 5))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n_e2e):
        e2e_step()
    e1.record()
    barrier()
    ms_e2e = e0.elapsed_time(e1) / n_e2e
    if world > 1:
        t = torch.tensor([ms_e2e], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_e2e = t.item()
    e2e = {"value": world * flops / (ms_e2e * 1e-3) / 1e12, "unit": "TFLOP/s", "ms_per_step": ms_e2e,
           "h2d_bytes_per_step": h_qkv.numel() * 2 + h_do.numel() * 2,
           "d2h_bytes_per_step": h_out.numel() * 2 + h_dqkv.numel() * 2}

    line = {
        "metric": METRIC, "value": value, "unit": "TFLOP/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": w["name"], "per_gpu": {"B": w["B"], "H": w["H"], "T": T, "D": D},
                   "global_batch": w["B"] * world, "causal": True, "api": "flash_attention_qkv_packed",
                   "chunks": [ncq, nckv], "parallelism": f"(batch,head) groups sharded over {world} GPU(s), no collective",
                   "l2": "inputs (805 MB/rank) larger than the 126 MB L2; no explicit flush"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline,
        "kernel_family": native.kernel_family(native.make_problem(qkv.detach()[0].unsqueeze(0), qkv.detach()[1].unsqueeze(0), True, 1.0)),
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        cpu_reference_step(1, T, D, threads)
        dt, fl = cpu_reference_step(2, T, D, threads)
        line["cpu_baseline"] = {"value": fl / dt / 1e12, "unit": "TFLOP/s", "cores": threads, "kind": "port",
                                "sample": f"2 of {G} (batch,head) groups at full T={T}, d={D}, fp32, torch CPU oracle port"}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_ring(args):
    """--workload ring: BASELINE.json configs[4] -- causal attention over ONE long sequence
    (T=131072, 32 heads, d=128, bf16) sharded zig-zag over the N GPUs, K/V exchanged by NCCL P2P
    around a ring with an fp32 LSE merge per hop (flash_attention_ipu_b200/ring.py).  Total work is
    fixed ("scaling": "strong").  Not the default bench line; the driver's contract line is run_ours."""
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist.init_process_group("nccl", device_id=dev, init_method=None if world > 1 else "tcp://127.0.0.1:29533",
                            rank=rank, world_size=world)
    from flash_attention_ipu_b200 import ring
    from flash_attention_ipu_b200._impl import native
    B, H, T, D = 1, 32, args.ring_seq, 128
    L2 = T // world
    torch.manual_seed(SEED + rank)
    q, k, v = (torch.randn(B, H, L2, D, device=dev).bfloat16().requires_grad_(True) for _ in range(3))
    dout = torch.randn(B, H, L2, D, device=dev).bfloat16()

    def step():
        q.grad = k.grad = v.grad = None
        ring.ring_attention(q, k, v).backward(dout)

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.3)
    launches0 = native.launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    barrier()
    m0 = sampler.mark()
    ev[0].record()
    for _ in range(args.steps):
        step()
    ev[1].record()
    barrier()
    m1 = sampler.mark()
    t = torch.tensor([ev[0].elapsed_time(ev[1])], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = t.item() / args.steps
    flops = attention_flops(B * H, T, D)
    line = {
        "metric": METRIC, "value": flops / (ms_step * 1e-3) / 1e12, "unit": "TFLOP/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": f"long_ctx_T{T}_H{H}_d{D}_bf16_causal_ring", "rows_per_gpu": L2,
                   "parallelism": f"sequence sharded zig-zag over {world} GPU(s); ring KV exchange by NCCL P2P "
                                  f"({2 * B * H * L2 * D * 2} B per hop per rank) + fp32 LSE merge",
                   "l2": "per-rank K/V shards larger than L2 for world<=8"},
        "clocks": sampler.summary(m0, m1), "gpu_launches": native.launch_count() - launches0,
    }
    if rank == 0:
        print(json.dumps(line), flush=True)
    dist.destroy_process_group()


OTHER_CONFIGS = {  # BASELINE.json configs[0], [1], [3]: reported with --workload, not the contract line
    "nanogpt": dict(name="nanogpt_attn_B64_T256_H6_d64_fp32_causal", B=64, H=6, T=256, D=64, dtype="fp32", causal=True),
    "gpt2": dict(name="gpt2_small_attn_B8_T1024_H12_d64_bf16_causal", B=8, H=12, T=1024, D=64, dtype="bf16", causal=True),
    "bert": dict(name="bert_large_attn_B32_T512_H16_d64_bf16_noncausal", B=32, H=16, T=512, D=64, dtype="bf16", causal=False),
}


def run_small(args):
    """--workload nanogpt|gpt2|bert: the small BASELINE.json configs on one GPU through the SDPA
    override (the call their models make).  These are launch/tail-bound (tens of microseconds of
    work); inputs fit in L2, so a 256 MB buffer is written between timed iterations to flush it."""
    import flash_attention_ipu.auto  # noqa: F401
    import torch.nn.functional as F
    from flash_attention_ipu_b200._impl import native
    w = OTHER_CONFIGS[args.workload]
    dev = torch.device("cuda", 0)
    dt = torch.float32 if w["dtype"] == "fp32" else torch.bfloat16
    torch.manual_seed(SEED)
    q, k, v = (torch.randn(w["B"], w["H"], w["T"], w["D"], device=dev, dtype=dt).requires_grad_(True) for _ in range(3))
    dout = torch.randn(w["B"], w["H"], w["T"], w["D"], device=dev, dtype=dt)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    sweep = [(0, 0)] if args.workload != "bert" else [(0, 0), (1, 1), (2, 2), (4, 2), (4, 4), (8, 8)]
    results = []
    for ncq, nckv in sweep:
        def step():
            q.grad = k.grad = v.grad = None
            if ncq == 0:
                out = F.scaled_dot_product_attention(q, k, v, is_causal=w["causal"])
            else:  # num_chunks sweep of config 4 through the packed API's keyword extension
                import flash_attention_ipu as fa
                qkv = torch.stack([q * w["D"] ** -0.5, k, v]).reshape(3, -1, w["T"], w["D"])
                out = fa.flash_attention_qkv_packed(qkv, ncq, nckv, causal=w["causal"]).view_as(q)
            out.backward(dout)
        for _ in range(args.warmup):
            step()
        torch.cuda.synchronize()
        launches0, total = native.launch_count(), 0.0
        for _ in range(args.steps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            step()
            e1.record()
            torch.cuda.synchronize()
            total += e0.elapsed_time(e1)
        ms = total / args.steps
        # the same step captured once into a CUDA graph and replayed: these shapes are launch-bound
        # (4 kernels of a few microseconds each behind Python dispatch)
        ms_graph = None
        try:
            gs = torch.cuda.Stream()
            gs.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(gs):
                step()
            torch.cuda.current_stream().wait_stream(gs)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                step()
            graph.replay()
            torch.cuda.synchronize()
            tg = 0.0
            for _ in range(args.steps):
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                graph.replay()
                e1.record()
                torch.cuda.synchronize()
                tg += e0.elapsed_time(e1)
            ms_graph = tg / args.steps
        except Exception as exc:  # capture is an optimisation of the harness, not of the product
            ms_graph = f"capture failed: {type(exc).__name__}"
        results.append({"chunks": [ncq, nckv], "ms_per_step": ms, "ms_per_step_cuda_graph": ms_graph,
                        "tflops": attention_flops(w["B"] * w["H"], w["T"], w["D"], causal=w["causal"]) / (ms * 1e-3) / 1e12,
                        "launches_per_step": (native.launch_count() - launches0) / args.steps})
    r0 = results[0]
    if isinstance(r0["ms_per_step_cuda_graph"], float):  # headline = the graph replay (what a training loop would run)
        r0 = dict(r0, ms_per_step=r0["ms_per_step_cuda_graph"],
                  tflops=r0["tflops"] * results[0]["ms_per_step"] / r0["ms_per_step_cuda_graph"])
    print(json.dumps({
        "metric": METRIC.replace("causal_", "" if w["causal"] else "noncausal_"), "value": r0["tflops"], "unit": "TFLOP/s",
        "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r0["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": w["dtype"], "data": "synthetic",
        "config": {"workload": w["name"], "api": "F.scaled_dot_product_attention (flash_attention_ipu.auto)",
                   "l2": "256 MB flush buffer written between timed iterations",
                   "launch": "fwd+bwd captured in one CUDA graph and replayed (eager time in chunk_sweep)",
                   "kernel_family": "tiled SIMT fp32 (FFMA2, csrc/fa_tiled_simt.cuh)" if w["dtype"] == "fp32" else "tcgen05"},
        "gpu_launches": int(r0["launches_per_step"] * args.steps), "chunk_sweep": results}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="llama", choices=["llama", "ring", "nanogpt", "gpt2", "bert"],
                    help="llama: configs[2], the contract line (default); ring: configs[4] long-context ring")
    ap.add_argument("--ring-seq", type=int, default=131072)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "ring":
        run_ring(args)
    elif args.workload in OTHER_CONFIGS:
        run_small(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
