#!/usr/bin/env python
"""Device time of a batched beam search (tuning aid, not a bench line).
    python tools/beam_time.py [images] [mem_len] [beam] [max_len] [precision]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from bench import START, END
from image_caption_b200 import CaptionDecodeEngine
from image_caption_b200.synthetic import random_decoder, random_memory

n_img = int(sys.argv[1]) if len(sys.argv) > 1 else 256
S = int(sys.argv[2]) if len(sys.argv) > 2 else 49
beam = int(sys.argv[3]) if len(sys.argv) > 3 else 5
max_len = int(sys.argv[4]) if len(sys.argv) > 4 else 50
prec = sys.argv[5] if len(sys.argv) > 5 else "bf16"
dev = torch.device("cuda:0")
dec = random_decoder(0, pe_rows=max(52, max_len)).to(dev)
mem = random_memory(n_img, S, seed=1234).to(dev)
eng = CaptionDecodeEngine(dec, precision=prec, max_images=n_img, max_group=beam, max_mem=S, max_len=max_len)
for _ in range(2):
    ids, lens, scores = eng.beam_search_from_memory(mem, START, END, max_len, beam_size=beam)
ts = []
for _ in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); ids, lens, scores = eng.beam_search_from_memory(mem, START, END, max_len, beam_size=beam); e1.record()
    torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
ms = sorted(ts)[2]
print(f"beam search {prec}: {n_img} images x beam {beam}, S={S}, max_len={max_len}: {ms:.2f} ms  "
      f"({n_img / ms * 1e3:.0f} captions/s, {1e3 * ms / (max_len - 1):.0f} us per step, mean length {float(lens.float().mean()):.1f})")
eng.close()
