#!/usr/bin/env python
"""Device time of capdec_prefill for a few (images, S) shapes, both GEMM routes (tuning aid, not a bench line).
    python tools/prefill_time.py [--shapes 1024x49,256x49,64x196,1024x100]"""
import argparse, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from image_caption_b200 import CaptionDecodeEngine
from image_caption_b200.synthetic import random_decoder, random_memory

ap = argparse.ArgumentParser()
ap.add_argument("--shapes", default="1024x49,256x49,64x196,1024x100,1024x196")
ap.add_argument("--reps", type=int, default=7)
a = ap.parse_args()
dev = torch.device("cuda:0")
dec = random_decoder(0, pe_rows=52).to(dev)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
for shp in a.shapes.split(","):
    n_img, S = map(int, shp.split("x"))
    mem = random_memory(n_img, S, seed=1234).to(dev)
    outs = {}
    for route in ("ws", "tile"):
        os.environ["CAPDEC_NO_WS_PREFILL"] = "0" if route == "ws" else "1"
        eng = CaptionDecodeEngine(dec, precision="bf16", max_images=n_img, max_group=1, max_mem=S, max_len=8)
        for i in range(3): eng.prefill(mem)
        ts = []
        for i in range(a.reps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); eng.prefill(mem); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = sorted(ts)[len(ts) // 2]
        outs[route] = torch.stack([eng.debug_cross_kv(l, w, n_img, S) for l in (0, 5) for w in (0, 1)])
        flops = 2.0 * n_img * S * 512 * 6144
        print(f"prefill {shp:9s} {route:4s}: {1e3 * ms:8.1f} us (convert + GEMM)  {flops / ms / 1e9:7.1f} TFLOP/s", flush=True)
        eng.close()
    print(f"   routes bit-equal: {bool(torch.equal(outs['ws'], outs['tile']))}  max|diff| {float((outs['ws'] - outs['tile']).abs().max()):.3e}", flush=True)
