#!/usr/bin/env python
"""Wall time per host-buffer rollout with and without capdec_prefetch_host (tuning aid)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from bench import WORKLOADS, START, END
from image_caption_b200 import CaptionDecodeEngine
from image_caption_b200.synthetic import random_decoder, random_memory
wl = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
desc, n_img, S, G, ns, max_len, _ = WORKLOADS[wl]
dev = torch.device("cuda:0")
dec = random_decoder(0, pe_rows=max(52, max_len)).to(dev)
mem = random_memory(n_img, S, seed=1234).pin_memory()
eng = CaptionDecodeEngine(dec, precision="bf16", max_images=n_img, max_group=G, max_mem=S, max_len=max_len)
run = lambda seed: eng.decode_from_memory(mem, START, END, max_len=max_len, group=G, n_sample=ns, seed=seed, trim=False)
for i in range(3): run(i)
for mode in ("plain", "prefetch", "plain", "prefetch"):
    torch.cuda.synchronize()
    ts = []
    tp = []
    td = []
    if mode == "prefetch": eng.prefetch_memory(mem)
    for i in range(8):
        t0 = time.perf_counter()
        if mode == "prefetch" and i + 1 < 8:
            eng.prefetch_memory(mem)
        t1 = time.perf_counter()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); run(10 + i); e1.record(); torch.cuda.synchronize()
        td.append(e0.elapsed_time(e1))
        ts.append(1e3 * (time.perf_counter() - t0)); tp.append(1e3 * (t1 - t0))
    print(f"{wl} {mode:9s}: per-step wall ms {' '.join(f'{t:6.2f}' for t in ts)}   prefetch call ms {max(tp):.3f}  device ms {' '.join(f'{t:6.2f}' for t in td)}", flush=True)
# the same overlap with torch alone: a side-stream H2D copy under a device-resident rollout
memd = mem.to(dev)
buf = torch.empty_like(memd)
side = torch.cuda.Stream()
rund = lambda seed: eng.decode_from_memory(memd, START, END, max_len=max_len, group=G, n_sample=ns, seed=seed, trim=False)
for mode in ("alone", "with_copy"):
    ts = []
    for i in range(4):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        if mode == "with_copy":
            with torch.cuda.stream(side): buf.copy_(mem, non_blocking=True)
        rund(20 + i); torch.cuda.synchronize()
        ts.append(1e3 * (time.perf_counter() - t0))
    print(f"{wl} device-resident rollout {mode:9s}: wall ms {' '.join(f'{t:6.2f}' for t in ts)}", flush=True)
t0 = time.perf_counter(); buf.copy_(mem, non_blocking=True); torch.cuda.synchronize(); print(f"H2D copy alone: {1e3 * (time.perf_counter() - t0):.2f} ms")
