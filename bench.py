#!/usr/bin/env python
"""bench.py - captions/s of batched caption decoding on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2|cfg3|cfg5|vit_greedy]
                    [--precision bf16|fp32] [--impl b200|reference]

A "step" is one whole rollout of the hot path over one batch: cross-K/V prefill plus
max_len-1 decode steps (BASELINE.json configs).  Default workload = configs[1]
(Grid 7x7 + Transformer greedy, 256 images, max_len 50) per GPU, weak scaling.

Prints ONE JSON line (rank 0).  `value` = captions/s with the encoder memory already resident in
HBM; `e2e` = the same through the host-buffer C-ABI call (pinned host memory in, token ids out,
copies inside the timed region); `roofline` = the dominant kernel (per-launch CUDA events from
the engine's profiling pass); `cpu_baseline` = the reference's no-KV-cache loop on the host cores.
`--impl reference` times only that CPU loop (the reference ships no GPU kernels of its own).
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
    # name: (description, images per GPU, S, group, n_sample, max_len)
    "cfg2": ("Grid(ResNet 7x7)+Transformer greedy decode, batch 256, max_len 50", 256, 49, 1, 0, 50),
    "cfg3": ("ViT+Transformer SCST rollout, 512 images x (5 samples + greedy), max_len 50", 512, 196, 6, 5, 50),
    "cfg5": ("Grid+Transformer greedy, 1024 images per GPU (8192 over 8), max_len 64", 1024, 49, 1, 0, 64),
    "vit_greedy": ("ViT+Transformer greedy decode, batch 512, max_len 50", 512, 196, 1, 0, 50),
}


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
            for name, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------- CPU reference arm
def cpu_reference(workload: str, steps: int, warmup: int, sample_images: int):
    """The reference's own algorithm (vit:307-323: full decoder over the whole prefix at every step,
    memory re-projected every call) on the host cores, on the first ``sample_images`` images of the
    workload.  /root/reference is absent on the GPU box, so the timed code is oracle/torch_port
    (same torch.nn modules, same loop)."""
    import torch
    from oracle import torch_port
    desc, n_img, S, G, ns, max_len = WORKLOADS[workload]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    dec = torch_port.build_decoder(0, max_len=max(52, max_len))
    mem = torch_port.synthetic_memory(sample_images, S)
    mem = mem.repeat_interleave(G, 0) if G > 1 else mem  # the reference decodes each rollout separately
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        ids = torch_port.greedy_recompute(dec, mem, START, END, max_len)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    assert ids.shape[1] == max_len
    caps = mem.shape[0]
    sec = sum(times) / len(times)
    return {"value": caps / sec, "unit": "captions/s", "cores": cores, "kind": "port",
            "sample": f"first {sample_images} of {n_img} images x {G} rollouts, all {max_len - 1} steps, "
                      f"reference no-KV-cache loop (oracle/torch_port.greedy_recompute), torch {torch.__version__} "
                      f"fp32, {cores} threads, mean of {len(times)}",
            "sec_per_sample": sec, "ms_per_decode_step": 1e3 * sec / (max_len - 1)}


def run_reference_arm(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    desc, n_img, S, G, ns, max_len = WORKLOADS[args.workload]
    r = cpu_reference(args.workload, args.steps, args.warmup, args.cpu_sample)
    line = {
        "impl": "reference", "metric": "captions/s", "value": r["value"], "unit": "captions/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * r["sec_per_sample"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {desc}", "images_per_gpu": n_img, "mem_len": S, "group": G,
                   "max_len": max_len, "note": "CPU reference arm: bounded sample, see cpu_baseline.sample"},
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": r["value"], "unit": "captions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "decode_step_us": 1e3 * r["ms_per_decode_step"],
    }
    print(json.dumps(line), flush=True)


def run_torch_gpu_arm(args) -> None:
    """Not part of the driver's contract: the reference's own loop (no KV cache, eager PyTorch fp32, oracle/torch_port)
    on ONE B200 over the whole workload -- SURVEY 8d's "real bar to beat".  Device time, CUDA events."""
    import torch
    from oracle import torch_port
    if int(os.environ.get("RANK", "0")) != 0:
        return
    desc, n_img, S, G, ns, max_len = WORKLOADS[args.workload]
    if args.images:
        n_img = args.images
    dev = torch.device("cuda", 0)
    dec = torch_port.build_decoder(0, max_len=max(52, max_len)).to(dev)
    mem = torch_port.synthetic_memory(n_img, S).to(dev)
    mem = mem.repeat_interleave(G, 0) if G > 1 else mem
    times = []
    for i in range(args.warmup + args.steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ids = torch_port.greedy_recompute(dec, mem, START, END, max_len)
        e1.record()
        torch.cuda.synchronize()
        if i >= args.warmup:
            times.append(e0.elapsed_time(e1))
    assert ids.shape[1] == max_len
    ms = sum(times) / len(times)
    print(json.dumps({"impl": "torch_eager_gpu", "metric": "captions/s", "value": mem.shape[0] / (ms * 1e-3), "unit": "captions/s",
                      "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
                      "dtype": "f32", "data": "synthetic", "decode_step_us": 1e3 * ms / (max_len - 1),
                      "config": {"workload": f"{args.workload}: {desc}", "sequences": mem.shape[0], "mem_len": S, "max_len": max_len,
                                 "note": f"reference no-KV-cache loop, eager torch {torch.__version__} fp32 on one B200, "
                                         "greedy for every sequence (sampling costs the same decoder calls)"}}), flush=True)


# --------------------------------------------------------------------------- B200 arm
def run_b200_arm(args) -> None:
    import torch
    import torch.distributed as dist
    from image_caption_b200 import CaptionDecodeEngine
    from image_caption_b200.dist import gather_rollout
    from image_caption_b200.synthetic import random_decoder, random_memory

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU path for the product arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    desc, n_img, S, G, ns, max_len = WORKLOADS[args.workload]
    if args.images:
        n_img = args.images
    if args.scaling == "strong":
        n_img = n_img // world
    M = n_img * G
    e = 2 if args.precision == "bf16" else 4
    T = max_len - 1

    dec = random_decoder(0, pe_rows=max(52, max_len)).to(dev)
    eng = CaptionDecodeEngine(dec, precision=args.precision, max_images=n_img, max_group=G, max_mem=S, max_len=max_len)
    # distinct images per rank (global image index seeds nothing but the slice)
    mem_host = random_memory(n_img, S, seed=1234 + rank).pin_memory()
    mem_dev = mem_host.to(dev)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    seq_base = rank * M
    counts = [n_img] * world

    def rollout(memory, seed):
        out = eng.decode_from_memory(memory, START, END, max_len=max_len, group=G, n_sample=ns, seed=seed,
                                     seq_id_base=seq_base, trim=False)
        if world > 1:  # the path's one collective: ids + log-probs for the reward step
            gather_rollout(out.ids, out.logp, counts, G)
        return out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- device-resident timing: per-step CUDA events, L2 flushed between steps (outside the events)
    for i in range(args.warmup):
        out = rollout(mem_dev, 2024 + i)
    assert out.steps == T, f"rollout stopped early ({out.steps} of {T} steps): random-init rows must not all end"
    barrier()
    launches0 = eng.launch_count()
    clocks = ClockSampler(local) if rank == 0 else None
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    for i in range(args.steps):
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
    ms_per_step = dev_ms / args.steps
    value = world * M * args.steps / (dev_ms * 1e-3)

    # ---- end to end through the host-buffer entry point (capdec_generate_host), wall clock
    for i in range(max(1, args.warmup // 2)):
        rollout(mem_host, 4000 + i)
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        oh = rollout(mem_host, 5000 + i)
    barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t.item())
    assert not oh.ids.is_cuda
    h2d = mem_host.numel() * 4
    d2h = M * max_len * 8 + (M * T * 4 if ns else 0)
    e2e = {"value": world * M * args.steps / e2e_s, "unit": "captions/s", "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": d2h, "ms_per_step": 1e3 * e2e_s / args.steps}

    line = {
        "metric": "captions/s", "value": value, "unit": "captions/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": args.scaling,
        "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
        "config": {"workload": f"{args.workload}: {desc}", "images_per_gpu": n_img, "mem_len": S, "group": G,
                   "n_sample": ns, "max_len": max_len, "sequences_per_gpu": M, "decode_steps": T,
                   "l2": "flushed between timed steps (512 MiB memset outside the event pairs)",
                   "weights": "random init, torch.manual_seed(0)", "collective": "all-gather ids+logp" if world > 1 else "none",
                   "engine": "v2 step kernels, 7 decode steps per CUDA graph, programmatic dependent launch"},
        "decode_step_us": 1e3 * ms_per_step / T, "e2e": e2e, "gpu_launches": int(launches), "clocks": clk,
    }

    if rank == 0:
        ab = algorithmic_bytes(n_img, S, G, max_len, e)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        hbm = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json)" if peaks else "fallback (B200_PROFILING.md)"
        line["rollout_roofline"] = {"bound": "hbm", "algorithmic_bytes": ab["rollout"],
                                    "achieved": ab["rollout"] / (ms_per_step * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s",
                                    "frac": ab["rollout"] / (ms_per_step * 1e-3) / 1e9 / hbm, "peak_source": peak_src}
        # per-kernel pass: same rollout, direct launches, one CUDA-event pair per launch
        eng.profile(True)
        flush.zero_()
        eng.decode_from_memory(mem_dev, START, END, max_len=max_len, group=G, n_sample=ns, seed=1, trim=False)
        rep = eng.profile_report()
        eng.profile(False)
        tot = sum(v["ms"] for v in rep.values())
        kernels = {k: {"launches": v["launches"], "us_per_launch": 1e3 * v["ms"] / v["launches"],
                       "share": v["ms"] / tot, "GBps": v["bytes"] / (v["ms"] * 1e-3) / 1e9 if v["ms"] > 0 else None}
                   for k, v in sorted(rep.items(), key=lambda kv: -kv[1]["ms"])}
        top = next(iter(kernels))
        tv = rep[top]
        traffic = None
        try:  # dram__bytes_read+write per launch of this kernel from the committed ncu --set full capture
            traffic = json.load(open(os.path.join(ROOT, "profiles", "r01_traffic.json"))).get(args.workload, {}).get(top)
        except OSError:
            pass
        line["roofline"] = {"kernel": top, "bound": "hbm", "achieved": tv["bytes"] / (tv["ms"] * 1e-3) / 1e9,
                            "peak": hbm, "unit": "GB/s", "frac": tv["bytes"] / (tv["ms"] * 1e-3) / 1e9 / hbm,
                            "traffic": traffic, "bytes_per_launch": tv["bytes"] / tv["launches"],
                            "us_per_launch": 1e3 * tv["ms"] / tv["launches"], "share_of_step": tv["ms"] / tot,
                            "peak_source": peak_src}
        line["kernels"] = kernels
        if world == 1 and not args.no_cpu:
            r = cpu_reference(args.workload, 1, 0, args.cpu_sample)
            line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line), flush=True)
    eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "torch_gpu"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--images", type=int, default=0, help="override images per GPU")
    ap.add_argument("--cpu-sample", type=int, default=8, help="images in the bounded CPU-baseline sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
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
