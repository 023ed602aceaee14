#!/usr/bin/env python
"""One warm rollout of a workload (for ncu captures).  python tools/one_rollout.py cfg2 [reps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from bench import WORKLOADS, START, END
from image_caption_b200 import CaptionDecodeEngine
from image_caption_b200.synthetic import random_decoder, random_memory
wl = sys.argv[1]; reps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
max_len_override = int(os.environ.get("MAXLEN", "0"))
desc, n_img, S, G, ns, max_len, _ = WORKLOADS[wl]
if max_len_override: max_len = max_len_override
dev = torch.device("cuda:0")
dec = random_decoder(0, pe_rows=max(52, max_len)).to(dev)
mem = random_memory(n_img, S, seed=1234).to(dev)
eng = CaptionDecodeEngine(dec, precision="bf16", max_images=n_img, max_group=G, max_mem=S, max_len=max_len,
                          use_graph=os.environ.get("NOGRAPH", "0") != "1")
for i in range(reps):
    out = eng.decode_from_memory(mem, START, END, max_len=max_len, group=G, n_sample=ns, seed=i, trim=False)
torch.cuda.synchronize()
print("steps", out.steps, "launches", eng.launch_count())
eng.close()
