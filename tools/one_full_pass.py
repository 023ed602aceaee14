#!/usr/bin/env python
"""Prefill + ONE full-sequence teacher-forced pass, bracketed by cudaProfilerStart/Stop, for ncu --profile-from-start off.
    python tools/one_full_pass.py cfg3 [--images K]"""
import argparse, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from bench import WORKLOADS, START, END
from image_caption_b200 import CaptionDecodeEngine
from image_caption_b200.synthetic import random_decoder, random_memory
ap = argparse.ArgumentParser()
ap.add_argument("workload")
ap.add_argument("--images", type=int, default=0)
a = ap.parse_args()
desc, n_img, S, G, ns, max_len, _ = WORKLOADS[a.workload]
if a.images: n_img = a.images
dev = torch.device("cuda:0")
dec = random_decoder(0, pe_rows=max(52, max_len)).to(dev)
mem = random_memory(n_img, S, seed=1234).to(dev)
eng = CaptionDecodeEngine(dec, precision="bf16", max_images=n_img, max_group=G, max_mem=S, max_len=max_len)
g = torch.Generator().manual_seed(5)
ids = torch.randint(4, 100, (n_img * G, max_len), generator=g).to(dev)
ids[:, 0] = START
eng.score(mem, ids, END, group=G, return_logits=False, mode="full")
torch.cuda.synchronize()
torch.cuda.profiler.start()
_, lp = eng.score(mem, ids, END, group=G, return_logits=False, mode="full")
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("rows", n_img * G * (max_len - 1), "mean logp", float(lp.mean()))
