#!/usr/bin/env python
"""bench.py - captions/s of batched caption decoding on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload auto|cfg2|cfg3|cfg4|cfg5|vit_greedy]
                    [--precision bf16|fp32|tf32] [--impl b200|reference|torch_gpu] [--no-sub] [--no-verify]

A "step" is one whole rollout of the hot path over one batch: cross-K/V prefill plus max_len-1 decode
steps (BASELINE.json configs).

Default workload (``auto``):
  * 1 GPU  -> cfg3, the largest single-GPU configuration (ViT + Transformer SCST rollout: 512 images x
    (5 samples + greedy), S = 196, max_len 50); cfg2, the cfg5 per-GPU shape, an fp32-mode run and the
    reference loop in eager PyTorch on the same B200 go into the same JSON line as ``sub`` records.
  * N GPUs -> cfg4 = the SAME 512-image global batch and seed sharded 512/N images per rank
    (``scaling: strong``), gathered by the path's one collective and verified bit-equal against an
    unsharded rollout on rank 0; cfg5 (1024 images per GPU) and cfg2 go in as weak-scaled ``sub`` records.

Prints ONE JSON line (rank 0).  ``value`` = captions/s with the encoder memory already resident in HBM;
``e2e`` = the same through the host-buffer entry points (pinned host memory in, token ids / log-probs back
on the host, copies inside the timed region); ``roofline`` = the dominant kernel (per-launch CUDA events
from the engine's profiling pass); ``cpu_baseline`` = the reference's no-KV-cache loop on the host cores.
``--impl reference`` times only that CPU loop on the same ``config`` (the reference ships no GPU kernels).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

START, END, PAD, VOCAB = 107, 108, 0, 109
D, L, F = 512, 6, 2048
P_STEP = 22_128_237      # parameters read by every decode step
P_PREFILL = 3_151_872    # cross K/V projection parameters (prefill only)

WORKLOADS = {
    # name: (description, images, S, group, n_sample, max_len, scaling); images = per GPU (weak) or global (strong)
    "cfg2": ("Grid(ResNet 7x7)+Transformer greedy decode, batch 256, max_len 50", 256, 49, 1, 0, 50, "weak"),
    "cfg3": ("ViT+Transformer SCST rollout, 512 images x (5 samples + greedy), max_len 50", 512, 196, 6, 5, 50, "weak"),
    "cfg4": ("ViT+Transformer SCST rollout, 512 images x (5 samples + greedy) sharded over the GPUs, max_len 50",
             512, 196, 6, 5, 50, "strong"),
    "cfg5": ("Grid+Transformer greedy, 1024 images per GPU (8192 over 8), max_len 64", 1024, 49, 1, 0, 64, "weak"),
    "vit_greedy": ("ViT+Transformer greedy decode, batch 512, max_len 50", 512, 196, 1, 0, 50, "weak"),
}
CPU_SAMPLE = {"cfg2": 8, "cfg3": 4, "cfg4": 4, "cfg5": 8, "vit_greedy": 8}  # images in the bounded CPU sample


def pick_workload(name: str, world: int) -> str:
    if name != "auto":
        return name
    return "cfg3" if world == 1 else "cfg4"


def shard_images(name: str, world: int, rank: int, images_override: int = 0):
    """(first image, images on this rank, images in the global batch)."""
    from image_caption_b200.dist import shard_bounds
    desc, n_img, S, G, ns, max_len, scaling = WORKLOADS[name]
    if images_override:
        n_img = images_override
    if scaling == "strong":
        lo, hi = shard_bounds(n_img, world, rank)
        return lo, hi - lo, n_img
    return rank * n_img, n_img, world * n_img


def workload_config(name: str, world: int, precision: str, images_override: int = 0) -> dict:
    """The ``config`` object: identical in the product arm and in the reference arm."""
    desc, n_img, S, G, ns, max_len, scaling = WORKLOADS[name]
    if images_override:
        n_img = images_override
    n_global = n_img if scaling == "strong" else n_img * world
    return {"workload": f"{name}: {desc}", "images_global": n_global, "images_per_gpu": n_global // world, "mem_len": S,
            "group": G, "n_sample": ns, "max_len": max_len, "sequences_global": n_global * G, "decode_steps": max_len - 1,
            "n_gpus": world, "scaling": scaling, "weights": "random init, torch.manual_seed(0)",
            "memory": "randn(images, S, 512), torch.Generator().manual_seed(1234)",
            # (the same object in the product arm and in the reference arm: statements about the GPU arm say so)
            "l2": "GPU arm: flushed between timed steps (512 MiB memset outside the event pairs)",
            "collective": ("GPU arm: all-gather ids+logp (NCCL), trimmed to the reference's global width" if world > 1 else "none"),
            "engine": "GPU arm: v2 step kernels (projection chains at <= 480 rows), 7 decode steps per CUDA graph, programmatic dependent launch"}


def algorithmic_bytes(n_img: int, S: int, G: int, max_len: int, e: int) -> dict:
    """SURVEY 8d.  e = bytes per stored weight / K / V element."""
    T = max_len - 1
    prefill = n_img * (S * D * 4 + 2 * L * S * D * e) + P_PREFILL * e
    steps = 0
    for n in range(1, T + 1):  # n = cached self positions read at this step
        steps += (P_STEP * e + n_img * 2 * L * S * D * e + n_img * G * (2 * L * n * D * e + 2 * L * D * e)
                  + n_img * G * (D * e + 8 + 4))
    return {"prefill": prefill, "steps": steps, "rollout": prefill + steps, "T": T}


class ClockSampler:
    """nvidia-smi clocks/throttle reasons while the timed region runs (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100", "-i", str(index)], stdout=subprocess.PIPE,
                                      stderr=subprocess.DEVNULL, text=True)
        except OSError:
            pass

    def stop(self) -> dict:
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            out, _ = self.p.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
            out, _ = self.p.communicate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------- CPU reference arm
def cpu_reference(name: str, steps: int, warmup: int, sample_images: int):
    """The reference's own algorithm (vit:307-323: full decoder over the whole prefix at every step, memory
    re-projected every call) on the host cores, on the first ``sample_images`` images of the workload's global batch,
    all rollouts of those images, all decode steps.  /root/reference is absent on the GPU box, so the timed code is
    oracle/torch_port (the same torch.nn modules in the same loop).  At least one un-timed warm-up run."""
    import torch
    from oracle import torch_port
    desc, n_img, S, G, ns, max_len, scaling = WORKLOADS[name]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    dec = torch_port.build_decoder(0, max_len=max(52, max_len))
    mem = torch_port.synthetic_memory(sample_images, S)
    mem = mem.repeat_interleave(G, 0) if G > 1 else mem  # the reference decodes each rollout separately
    times = []
    for i in range(max(1, warmup) + steps):
        t0 = time.perf_counter()
        ids = torch_port.greedy_recompute(dec, mem, START, END, max_len)
        dt = time.perf_counter() - t0
        if i >= max(1, warmup):
            times.append(dt)
    assert ids.shape[1] == max_len
    caps = mem.shape[0]
    sec = sum(times) / len(times)
    return {"value": caps / sec, "unit": "captions/s", "cores": cores, "kind": "port",
            "sample": f"first {sample_images} images x {G} rollouts of the workload's batch, all {max_len - 1} decode steps, "
                      f"reference no-KV-cache loop (oracle/torch_port.greedy_recompute), torch {torch.__version__} fp32, "
                      f"{cores} threads, {max(1, warmup)} warm-up + mean of {len(times)} runs",
            "sec_per_sample": sec, "ms_per_decode_step": 1e3 * sec / (max_len - 1)}


def run_reference_arm(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    name = pick_workload(args.workload, world)
    desc, n_img, S, G, ns, max_len, scaling = WORKLOADS[name]
    r = cpu_reference(name, args.steps, args.warmup, args.cpu_sample or CPU_SAMPLE[name])
    line = {
        "impl": "reference", "metric": "captions/s", "value": r["value"], "unit": "captions/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * r["sec_per_sample"], "higher_is_better": True, "scaling": scaling,
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(name, world, args.precision, args.images),
        "sample": r["sample"],
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": r["value"], "unit": "captions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "decode_step_us": 1e3 * r["ms_per_decode_step"],
    }
    print(json.dumps(line), flush=True)


def torch_eager_gpu(name: str, steps: int, warmup: int, local: int, images_override: int = 0) -> dict:
    """The reference's own loop (no KV cache, eager PyTorch fp32, oracle/torch_port) on ONE B200 over the whole
    workload -- SURVEY 8d's "real bar to beat".  Device time (CUDA events), clocks sampled during the timed runs."""
    import torch
    from oracle import torch_port
    desc, n_img, S, G, ns, max_len, scaling = WORKLOADS[name]
    if images_override:
        n_img = images_override
    dev = torch.device("cuda", local)
    dec = torch_port.build_decoder(0, max_len=max(52, max_len)).to(dev)
    mem = torch_port.synthetic_memory(n_img, S).to(dev)
    mem = mem.repeat_interleave(G, 0) if G > 1 else mem
    for _ in range(warmup):
        torch_port.greedy_recompute(dec, mem, START, END, max_len)
    torch.cuda.synchronize(dev)
    clocks = ClockSampler(local)
    times = []
    for _ in range(steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ids = torch_port.greedy_recompute(dec, mem, START, END, max_len)
        e1.record()
        torch.cuda.synchronize(dev)
        times.append(e0.elapsed_time(e1))
    clk = clocks.stop()
    assert ids.shape[1] == max_len
    ms = sum(times) / len(times)
    del dec, mem
    torch.cuda.empty_cache()
    return {"impl": "torch_eager_gpu", "metric": "captions/s", "value": n_img * G / (ms * 1e-3), "unit": "captions/s",
            "n_gpus": 1, "steps": steps, "warmup": warmup, "ms_per_step": ms, "dtype": "f32",
            "decode_step_us": 1e3 * ms / (max_len - 1), "clocks": clk, "sequences": n_img * G,
            "note": f"reference no-KV-cache loop (oracle/torch_port.greedy_recompute), eager torch {torch.__version__} fp32 on "
                    "one B200, greedy for every sequence (sampling costs the same decoder calls)"}


def teacher_forced(name: str, local: int, reps: int = 3) -> dict:
    """SURVEY 8f-4 sub-record: log-probs of GIVEN captions (the SCST gradient pass's forward, scst:183-191 / vit:216-255)
    for the whole workload batch -- the engine's step replay, its full-sequence pass (capdec_forward_full), and the
    reference's single causal decoder call in eager PyTorch fp32 on the same B200.  Device time (CUDA events, L2 flushed
    between runs), prefill included in both engine figures."""
    import torch
    from image_caption_b200 import CaptionDecodeEngine
    from image_caption_b200.synthetic import random_decoder, random_memory
    from oracle import torch_port
    desc, n_img, S, G, ns, max_len, scaling = WORKLOADS[name]
    dev = torch.device("cuda", local)
    dec = random_decoder(0, pe_rows=max(52, max_len)).to(dev)
    mem = random_memory(n_img, S, seed=1234).to(dev)
    eng = CaptionDecodeEngine(dec, precision="bf16", max_images=n_img, max_group=G, max_mem=S, max_len=max_len)
    ids = eng.decode_from_memory(mem, START, END, max_len=max_len, group=G, n_sample=ns, seed=1, trim=False).ids.clone()
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    M, T = n_img * G, max_len - 1

    def timed(fn, n):
        ts = []
        for _ in range(n):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); out = fn(); e1.record()
            torch.cuda.synchronize(dev)
            ts.append(e0.elapsed_time(e1))
        return sorted(ts)[len(ts) // 2], out

    rec = {"metric": "captions/s", "unit": "captions/s", "dtype": "bf16", "sequences": M, "positions": T,
           "what": "log p(ids[:, k+1] | ids[:, :k+1]) for every position of every sequence of the workload batch"}
    outs = {}
    for mode in ("steps", "full"):
        for _ in range(2):
            eng.score(mem, ids, END, group=G, return_logits=False, mode=mode)
        ms, out = timed(lambda: eng.score(mem, ids, END, group=G, return_logits=False, mode=mode), reps)
        outs[mode] = out[1]
        rec[mode] = {"ms": ms, "value": M / (ms * 1e-3)}
    rec["value"] = rec["full"]["value"]
    rec["speedup_full_over_steps"] = rec["steps"]["ms"] / rec["full"]["ms"]
    rec["max_abs_logp_diff_full_vs_steps"] = float((outs["full"] - outs["steps"]).abs().max())
    eng.profile(True)
    eng.score(mem, ids, END, group=G, return_logits=False, mode="full")
    rep = eng.profile_report()
    eng.profile(False)
    tot = sum(v["ms"] for v in rep.values())
    rec["kernels"] = {k: {"launches": v["launches"], "us_per_launch": 1e3 * v["ms"] / v["launches"], "share": v["ms"] / tot}
                      for k, v in sorted(rep.items(), key=lambda kv: -kv[1]["ms"])}
    # tensor roofline of the projections: 2 * rows * (1536 + 512 + 512 + 2048 + 2048) * 512 ... per layer
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    peak = float(peaks.get("bf16_tflops_sustained", 1355.1))
    flops = 2.0 * M * T * 512 * (1536 + 512 + 512 + 512 + 2048 + 2048) * 6
    gemm_ms = sum(v["ms"] for k, v in rep.items() if k.startswith("full_gemm") and k != "full_gemm_vocab")
    rec["roofline"] = {"kernel": "full_gemm_* (the six projections of every layer)", "bound": "tensor",
                       "achieved": flops / (gemm_ms * 1e-3) / 1e12, "peak": peak, "unit": "TFLOP/s",
                       "frac": flops / (gemm_ms * 1e-3) / 1e12 / peak, "traffic": None,
                       "peak_source": "measured sustained bf16 (MEASURED_PEAKS.json)" if peaks else "fallback (B200_PROFILING.md)"}
    eng.close()
    del eng
    # the reference's own call: one causal decoder pass in eager fp32 (oracle/torch_port), chunked over sequences to bound memory
    pdec = torch_port.build_decoder(0, max_len=max(52, max_len)).to(dev)
    memr = mem.repeat_interleave(G, 0) if G > 1 else mem

    def eager():
        with torch.no_grad():
            outs = []
            for lo in range(0, M, 512):
                lg = torch_port.teacher_forced_logits(pdec, ids[lo:lo + 512], memr[lo:lo + 512])
                outs.append(torch.log_softmax(lg, -1).gather(2, ids[lo:lo + 512, 1:, None]).squeeze(2))
            return torch.cat(outs)
    eager()
    ms, _ = timed(eager, 2)
    rec["torch_eager_gpu_fp32"] = {"ms": ms, "value": M / (ms * 1e-3),
                                   "note": "reference decoder, one causal pass per 512 sequences, eager torch fp32 on the same B200"}
    del pdec, memr, mem, flush
    torch.cuda.empty_cache()
    return rec


def run_torch_gpu_arm(args) -> None:
    if int(os.environ.get("RANK", "0")) != 0:
        return
    name = pick_workload(args.workload, 1)
    r = torch_eager_gpu(name, args.steps, max(1, args.warmup), 0, args.images)
    r.update(higher_is_better=True, data="synthetic", config=workload_config(name, 1, "fp32", args.images))
    print(json.dumps(r), flush=True)


# --------------------------------------------------------------------------- B200 arm
def measure(name: str, precision: str, steps: int, warmup: int, world: int, rank: int, local: int, images_override: int = 0,
            profile: bool = True, verify: bool = False) -> dict:
    """Time one workload on this process group; returns the record (complete on rank 0)."""
    import torch
    import torch.distributed as dist
    from image_caption_b200 import CaptionDecodeEngine
    from image_caption_b200.dist import HostFeeder, sharded_rollout
    from image_caption_b200.synthetic import random_decoder, random_memory

    dev = torch.device("cuda", local)
    desc, _, S, G, ns, max_len, scaling = WORKLOADS[name]
    lo, n_img, n_global = shard_images(name, world, rank, images_override)
    M = n_img * G
    e = 2 if precision == "bf16" else 4
    T = max_len - 1

    dec = random_decoder(0, pe_rows=max(52, max_len)).to(dev)
    eng = CaptionDecodeEngine(dec, precision=precision, max_images=n_img, max_group=G, max_mem=S, max_len=max_len)
    # ONE global batch (seed 1234); every rank takes its contiguous block of images, so the sharded job decodes the
    # same images with the same Philox sequence ids (seq_id_base = first image * group) at every N
    mem_global = random_memory(n_global, S, seed=1234)
    mem_host = mem_global[lo:lo + n_img].clone().pin_memory()
    del mem_global
    mem_dev = mem_host.to(dev)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def local_decode(memory, seed):
        return eng.decode_from_memory(memory, START, END, max_len=max_len, group=G, n_sample=ns, seed=seed,
                                      seq_id_base=lo * G, trim=False, early_stop=(world == 1))

    def rollout(memory, seed):
        """Device-resident input; N > 1: the path's one collective (ids + log-probs for the reward step)."""
        if world == 1:
            out = local_decode(memory, seed)
            return out.ids, out.logp, out.steps

        def shard(lo_, hi_, base):
            assert lo_ == lo and hi_ == lo + n_img and base == lo * G
            o = local_decode(memory, seed)
            return o.ids, o.logp
        ids, logp = sharded_rollout(shard, n_global, START, END, max_len, group=G, n_sample=ns, seed=seed)
        return ids, logp, ids.shape[1] - 1

    def rollout_host(seed):
        """Host buffers in, host results out.  1 GPU: the C-ABI host entry point (capdec_generate_host).  N GPUs: H2D of
        this rank's shard, decode, gather ON THE DEVICE, then one D2H of the gathered result."""
        if world == 1:
            out = local_decode(mem_host, seed)
            return out.ids, out.logp
        ids, logp, _ = rollout(feeder.take(), seed)
        ids_h = ids.cpu()
        logp_h = logp.cpu() if logp is not None else None
        return ids_h, logp_h

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def prefetch_host():
        """Start the H2D copy of one step's input: under the C ABI (1 GPU) or on a side stream (dist.HostFeeder)."""
        if world == 1:
            eng.prefetch_memory(mem_host)
        else:
            feeder.prefetch(mem_host)

    # ---- device-resident timing: per-step CUDA events, L2 flushed between steps (outside the events)
    for i in range(warmup):
        ids, logp, n_steps = rollout(mem_dev, 2024 + i)
    assert n_steps == T, f"rollout stopped early ({n_steps} of {T} steps): random-init rows must not all end"
    barrier()
    launches0 = eng.launch_count()
    clocks = ClockSampler(local) if rank == 0 else None
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for i in range(steps):
        flush.zero_()
        ev[i][0].record()
        rollout(mem_dev, 3000 + i)
        ev[i][1].record()
    barrier()
    clk = clocks.stop() if clocks else None
    launches = eng.launch_count() - launches0
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    t = torch.tensor([dev_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms = float(t.item())
    ms_per_step = dev_ms / steps
    value = n_global * G * steps / (dev_ms * 1e-3)

    # ---- end to end through the host-buffer entry points, wall clock
    feeder = HostFeeder(dev) if world > 1 else None
    for i in range(max(1, warmup // 2)):
        prefetch_host()
        rollout_host(4000 + i)
    barrier()
    t0 = time.perf_counter()
    prefetch_host()  # step 0's H2D copy (exposed)
    for i in range(steps):
        if i + 1 < steps:
            prefetch_host()  # step i+1's H2D copy starts now and runs underneath step i's decode
        ids_h, logp_h = rollout_host(5000 + i)
    barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t.item())
    assert not ids_h.is_cuda
    h2d = mem_host.numel() * 4
    d2h = ids_h.numel() * 8 + (logp_h.numel() * 4 if logp_h is not None else 0)
    e2e = {"value": n_global * G * steps / e2e_s, "unit": "captions/s", "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": d2h, "ms_per_step": 1e3 * e2e_s / steps,
           "path": "capdec_generate_host (pinned host memory in, ids/log-probs out; every step copies its memory H2D inside the timed region: capdec_prefetch_host starts step i+1's copy before step i's call, so it runs under step i's decode; step 0's copy is exposed)"
                   if world == 1 else "per rank: H2D of the shard (dist.HostFeeder: step i+1's copy on a side stream under step i's decode, every copy inside the timed region), decode, all-gather on the device, one D2H of the gathered result"}

    cfg = workload_config(name, world, precision, images_override)
    rec = {"metric": "captions/s", "value": value, "unit": "captions/s", "n_gpus": world, "steps": steps,
           "warmup": warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": scaling,
           "vs_baseline": None, "dtype": precision, "data": "synthetic", "config": cfg,
           "decode_step_us": 1e3 * ms_per_step / T, "e2e": e2e, "gpu_launches": int(launches), "clocks": clk}

    # ---- sharded result vs unsharded result (same global batch, same seed): rank 0 decodes everything alone.
    # (1) as measured (fastest kernel route per batch size): bit-equal when the shard and the whole batch take the same
    #     route, else the suite's parity criterion applied across the sharding: the unsharded engine scores the gathered
    #     ids teacher-forced; sampled log-probs within the bf16 tolerance, greedy tokens = its argmax outside the tie tolerance;
    # (2) with CAPDEC_FLAG_BATCH_INVARIANT (one route for every batch size): bit-equal at every N.
    if verify and world > 1:
        tol = {"bf16": 4e-2, "tf32": 2e-2}.get(precision, 2e-4)
        ok = {}
        for mode in ("as_measured", "batch_invariant"):
            inv = mode == "batch_invariant"
            if inv and precision != "bf16":
                continue
            eng_s = eng if not inv else CaptionDecodeEngine(dec, precision=precision, max_images=n_img, max_group=G, max_mem=S,
                                                            max_len=max_len, batch_invariant=True)

            def shard(lo_, hi_, base, eng_s=eng_s):
                o = eng_s.decode_from_memory(mem_dev, START, END, max_len=max_len, group=G, n_sample=ns, seed=777,
                                             seq_id_base=lo * G, trim=False, early_stop=False)
                return o.ids, o.logp
            ids_s, logp_s = sharded_rollout(shard, n_global, START, END, max_len, group=G, n_sample=ns, seed=777)
            if inv:
                eng_s.close()
            if rank == 0:
                eng1 = CaptionDecodeEngine(dec, precision=precision, max_images=n_global, max_group=G, max_mem=S, max_len=max_len,
                                           batch_invariant=inv)
                mem_all = random_memory(n_global, S, seed=1234).to(dev)
                full = eng1.decode_from_memory(mem_all, START, END, max_len=max_len, group=G, n_sample=ns, seed=777, seq_id_base=0,
                                               trim=False, early_stop=False)
                w = ids_s.shape[1]
                r = {"ids_bit_equal": bool(torch.equal(full.ids[:, :w], ids_s)),
                     "logp_bit_equal": bool(torch.equal(full.logp[:, :w - 1], logp_s)) if logp_s is not None else None,
                     "rows_with_a_different_token": int((full.ids[:, :w] != ids_s).any(1).sum()), "width": int(w)}
                if not inv:
                    pad_ids = torch.full((ids_s.shape[0], max_len), PAD, dtype=torch.int64, device=dev)
                    pad_ids[:, :w] = ids_s
                    logits, lp = eng1.score(mem_all, pad_ids, END, group=G)
                    logits, lp = logits[:, :w - 1], lp[:, :w - 1]
                    is_s = (torch.arange(ids_s.shape[0], device=dev) % G) < ns
                    top2 = logits.topk(2, -1).values
                    gap = top2[..., 0] - top2[..., 1]
                    bad = (ids_s[:, 1:] != logits.argmax(-1)) & (gap > tol) & ~is_s[:, None]
                    r["greedy_tokens_off_argmax_outside_tolerance"] = int(bad.sum())
                    if logp_s is not None and ns:
                        r["sampled_logp_max_abs_diff"] = float((lp[is_s] - logp_s[is_s]).abs().max())
                        r["within_tolerance"] = bool(r["sampled_logp_max_abs_diff"] <= tol and int(bad.sum()) == 0)
                    else:
                        r["within_tolerance"] = bool(int(bad.sum()) == 0)
                    r["tolerance"] = tol
                ok[mode] = r
                eng1.close()
                del mem_all
            barrier()
        ok["what"] = (f"gathered {world}-rank rollout vs ONE unsharded rollout of the same {n_global} images on rank 0, same seed; "
                      "as_measured = fastest kernel route per batch size (bit-equality only when shard and whole batch share a route), "
                      "batch_invariant = CAPDEC_FLAG_BATCH_INVARIANT engines on both sides")
        rec["verify"] = ok if rank == 0 else None
        barrier()

    if rank == 0:
        ab = algorithmic_bytes(n_img, S, G, max_len, e)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        hbm = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json)" if peaks else "fallback (B200_PROFILING.md)"
        rec["rollout_roofline"] = {"bound": "hbm", "algorithmic_bytes": ab["rollout"],
                                   "achieved": ab["rollout"] / (ms_per_step * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s",
                                   "frac": ab["rollout"] / (ms_per_step * 1e-3) / 1e9 / hbm, "peak_source": peak_src,
                                   "note": "per GPU: this rank's shard"}
        if profile:
            # per-kernel pass: same rollout, direct launches, one CUDA-event pair per launch (each pair carries ~5 us of
            # launch/event floor: the SHARES are meaningful, the absolute per-launch times are upper bounds; the ncu
            # gpu__time_duration list of the same command is committed under profiles/)
            eng.profile(True)
            flush.zero_()
            eng.decode_from_memory(mem_dev, START, END, max_len=max_len, group=G, n_sample=ns, seed=1, trim=False)
            rep = eng.profile_report()
            eng.profile(False)
            tot = sum(v["ms"] for v in rep.values())
            kernels = {k: {"launches": v["launches"], "us_per_launch": 1e3 * v["ms"] / v["launches"],
                           "share": v["ms"] / tot, "GBps": v["bytes"] / (v["ms"] * 1e-3) / 1e9 if v["ms"] > 0 else None}
                       for k, v in sorted(rep.items(), key=lambda kv: -kv[1]["ms"])}
            ncu = {}
            try:  # per-kernel gpu__time_duration (us) and dram bytes per launch from the committed ncu captures
                ncu = json.load(open(os.path.join(ROOT, "profiles", "ncu_kernels.json"))).get(name, {})
            except (OSError, ValueError):
                pass
            for k in kernels:
                if k in ncu:
                    kernels[k]["ncu_us_per_launch"] = ncu[k].get("us")
            top = next(iter(kernels))
            tv = rep[top]
            rec["roofline"] = {"kernel": top, "bound": "hbm", "achieved": tv["bytes"] / (tv["ms"] * 1e-3) / 1e9,
                               "peak": hbm, "unit": "GB/s", "frac": tv["bytes"] / (tv["ms"] * 1e-3) / 1e9 / hbm,
                               "traffic": ncu.get(top, {}).get("dram_bytes"), "bytes_per_launch": tv["bytes"] / tv["launches"],
                               "us_per_launch": 1e3 * tv["ms"] / tv["launches"], "share_of_step": tv["ms"] / tot,
                               "ncu_us_per_launch": ncu.get(top, {}).get("us"), "peak_source": peak_src}
            rec["kernels"] = kernels
    eng.close()
    del eng, flush, mem_dev
    torch.cuda.empty_cache()
    return rec


def run_b200_arm(args) -> None:
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU path for the product arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    name = pick_workload(args.workload, world)
    line = measure(name, args.precision, args.steps, args.warmup, world, rank, local, args.images,
                   verify=not args.no_verify)

    if not args.no_sub and args.workload == "auto":
        sub = {}
        sub_steps = max(3, args.steps // 2)
        for sname in (("cfg2", "cfg5") if world == 1 else ("cfg5", "cfg2")):
            r = measure(sname, args.precision, sub_steps, args.warmup, world, rank, local)
            sub[sname if sname != "cfg5" else "cfg5_per_gpu_1024"] = r
        if world == 1:
            r = measure(name, "fp32", 3, 3, world, rank, local, profile=False)
            sub[f"{name}_fp32"] = r
            r = measure(name, "tf32", 3, 3, world, rank, local, profile=False)  # fp32 storage, tcgen05 kind::tf32 projections
            sub[f"{name}_tf32"] = r
            if rank == 0:
                sub[f"torch_eager_gpu_{name}"] = torch_eager_gpu(name, 1, 1, local)
                sub[f"teacher_forced_{name}"] = teacher_forced(name, local)
        line["sub"] = sub
    if rank == 0:
        if world == 1 and not args.no_cpu:
            r = cpu_reference(name, 3, 1, args.cpu_sample or CPU_SAMPLE[name])
            line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "torch_gpu"])
    ap.add_argument("--workload", default="auto", choices=["auto"] + sorted(WORKLOADS))
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32", "tf32"])
    ap.add_argument("--images", type=int, default=0, help="override the workload's image count")
    ap.add_argument("--cpu-sample", type=int, default=0, help="images in the bounded CPU-baseline sample (0 = per workload)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-sub", action="store_true", help="skip the sub-records of the default workload")
    ap.add_argument("--no-verify", action="store_true", help="N > 1: skip the sharded == unsharded check")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    if args.impl == "reference":
        run_reference_arm(args)
    elif args.impl == "torch_gpu":
        run_torch_gpu_arm(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()
