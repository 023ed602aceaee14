#!/usr/bin/env python
"""Encoder hand-off: device time of `projection (torch) + capdec_prefill` against `capdec_prefill_tokens`
for a ViT-shaped batch (tuning aid).   python tools/handoff_time.py [--images 512]"""
import argparse, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_caption_b200 import CaptionDecodeEngine
from image_caption_b200.synthetic import random_decoder

ap = argparse.ArgumentParser()
ap.add_argument("--images", type=int, default=512)
ap.add_argument("--reps", type=int, default=10)
a = ap.parse_args()
dev = torch.device("cuda:0")
dec = random_decoder(0, pe_rows=52).to(dev)
eng = CaptionDecodeEngine(dec, precision="bf16", max_images=a.images, max_group=1, max_mem=196, max_len=50)
g = torch.Generator(device=dev).manual_seed(1)
tok = torch.randn(a.images, 197, 768, device=dev, generator=g)
proj = torch.nn.Linear(768, 512).to(dev)

def timed(fn):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / a.reps

with torch.no_grad():
    t_ref = timed(lambda: eng.prefill(proj(tok[:, 1:, :])))
    t_tf32 = None
    torch.backends.cuda.matmul.allow_tf32 = True
    t_tf32 = timed(lambda: eng.prefill(proj(tok[:, 1:, :])))
    torch.backends.cuda.matmul.allow_tf32 = False
    t_new = timed(lambda: eng.prefill_tokens(tok, proj, first_token=1))
print(f"{a.images} images x 196 tokens: torch projection (fp32) + prefill {t_ref:.3f} ms; with TF32 matmul {t_tf32:.3f} ms; "
      f"prefill_tokens (projection folded) {t_new:.3f} ms")
eng.close()
