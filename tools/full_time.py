#!/usr/bin/env python
"""Teacher-forced scoring: step replay vs the full-sequence pass (tuning aid).  python tools/full_time.py cfg3 [--images K]"""
import argparse, os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from bench import WORKLOADS, START, END
from image_caption_b200 import CaptionDecodeEngine
from image_caption_b200.synthetic import random_decoder, random_memory

ap = argparse.ArgumentParser()
ap.add_argument("workload")
ap.add_argument("--images", type=int, default=0)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--profile", action="store_true")
a = ap.parse_args()
desc, n_img, S, G, ns, max_len, _ = WORKLOADS[a.workload]
if a.images: n_img = a.images
dev = torch.device("cuda:0")
dec = random_decoder(0, pe_rows=max(52, max_len)).to(dev)
mem = random_memory(n_img, S, seed=1234).to(dev)
eng = CaptionDecodeEngine(dec, precision="bf16", max_images=n_img, max_group=G, max_mem=S, max_len=max_len)
out = eng.decode_from_memory(mem, START, END, max_len=max_len, group=G, n_sample=ns, seed=1, trim=False)
ids = out.ids.clone()
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
res = {}
for mode in ("steps", "full"):
    for i in range(2): eng.score(mem, ids, END, group=G, return_logits=False, mode=mode)
    ts = []
    for i in range(a.reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); r = eng.score(mem, ids, END, group=G, return_logits=False, mode=mode); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    res[mode] = (sorted(ts)[len(ts) // 2], r[1])
    print(f"{a.workload} score mode={mode:5s}: {res[mode][0]:9.3f} ms  {n_img * G / res[mode][0] * 1e3:9.0f} captions/s (incl. prefill)", flush=True)
print("max |logp full - steps|:", float((res['full'][1] - res['steps'][1]).abs().max()))
if a.profile:
    eng.profile(True)
    eng.score(mem, ids, END, group=G, return_logits=False, mode="full")
    rep = eng.profile_report()
    eng.profile(False)
    tot = sum(v["ms"] for v in rep.values())
    for k, v in sorted(rep.items(), key=lambda kv: -kv[1]["ms"]):
        print(f"   {k:26s} n={v['launches']:4d} us/launch={1e3 * v['ms'] / v['launches']:9.1f} share={v['ms'] / tot:.3f} total_ms={v['ms']:.3f}")
