#!/usr/bin/env python
"""Device time per decode step for one workload and a few engine variants (tuning aid, not a bench line).
    python tools/step_time.py cfg2 [--variants v2,v2_nopdl,v1] [--bnb 32,64]"""
import argparse, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from bench import WORKLOADS, START, END
from image_caption_b200 import CaptionDecodeEngine
from image_caption_b200.synthetic import random_decoder, random_memory

ap = argparse.ArgumentParser()
ap.add_argument("workload")
ap.add_argument("--variants", default="v2,v1")
ap.add_argument("--bnb", default="")
ap.add_argument("--stages", default="")
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--images", type=int, default=0)
a = ap.parse_args()
desc, n_img, S, G, ns, max_len, _ = WORKLOADS[a.workload]
if a.images: n_img = a.images
dev = torch.device("cuda:0")
dec = random_decoder(0, pe_rows=max(52, max_len)).to(dev)
mem = random_memory(n_img, S, seed=1234).to(dev)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
for var in a.variants.split(","):
    for bnb in (a.bnb.split(",") if a.bnb else [""]):
        for stg in (a.stages.split(",") if a.stages else [""]):
            os.environ.pop("CAPDEC_BNB", None); os.environ.pop("CAPDEC_STAGES", None)
            if bnb: os.environ["CAPDEC_BNB"] = bnb
            if stg: os.environ["CAPDEC_STAGES"] = stg
            kw = dict(step_impl=var[:2], pdl=(var != "v2_nopdl"), use_graph=(var != "v2_nograph"))
            eng = CaptionDecodeEngine(dec, precision="bf16", max_images=n_img, max_group=G, max_mem=S, max_len=max_len, **kw)
            run = lambda seed: eng.decode_from_memory(mem, START, END, max_len=max_len, group=G, n_sample=ns, seed=seed, trim=False)
            for i in range(3): run(i)
            ts = []
            for i in range(a.reps):
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); run(10 + i); e1.record(); torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            ms = sorted(ts)[len(ts) // 2]
            print(f"{a.workload} {var:10s} bnb={bnb or 'auto':4s} stages={stg or 'auto':4s}: rollout {ms:8.3f} ms  "
                  f"step {1e3 * ms / (max_len - 1):8.1f} us  {n_img * G / ms * 1e3:9.0f} captions/s", flush=True)
            eng.close()
