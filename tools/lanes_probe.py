#!/usr/bin/env python
"""Probe: one rollout split into `lanes` independent sub-batches of images, each on its own engine handle and
CUDA stream, so that the tensor-bound GEMMs of one lane overlap the HBM-bound attention of another (tuning aid).
    python tools/lanes_probe.py cfg3 --lanes 1,2,3,4 [--prio]"""
import argparse, ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from bench import WORKLOADS, START, END
from image_caption_b200 import CaptionDecodeEngine, _lib
from image_caption_b200.synthetic import random_decoder, random_memory

ap = argparse.ArgumentParser()
ap.add_argument("workload")
ap.add_argument("--lanes", default="1,2,3,4")
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--prio", action="store_true", help="alternate stream priorities between lanes")
a = ap.parse_args()
desc, n_img, S, G, ns, max_len, _ = WORKLOADS[a.workload]
dev = torch.device("cuda:0")
dec = random_decoder(0, pe_rows=max(52, max_len)).to(dev)
mem = random_memory(n_img, S, seed=1234).to(dev)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
lib = _lib.load()
ref_ids = None
for lanes in [int(x) for x in a.lanes.split(",")]:
    bounds = [(n_img * i) // lanes for i in range(lanes + 1)]
    engs, streams, outs = [], [], []
    for i in range(lanes):
        cnt = bounds[i + 1] - bounds[i]
        e = CaptionDecodeEngine(dec, precision="bf16", max_images=cnt, max_group=G, max_mem=S, max_len=max_len)
        e.sync_weights()
        engs.append(e)
        streams.append(torch.cuda.Stream(device=dev, priority=(-1 if (a.prio and i % 2 == 0) else 0)))
        outs.append((torch.empty((cnt * G, max_len), dtype=torch.int64, device=dev),
                     torch.empty((cnt * G, max_len - 1), dtype=torch.float32, device=dev)))
    torch.cuda.synchronize()

    def run(seed):
        main = torch.cuda.current_stream(dev)
        ev0 = torch.cuda.Event(); ev0.record(main)
        for i, e in enumerate(engs):
            s = streams[i]
            s.wait_event(ev0)
            lo, hi = bounds[i], bounds[i + 1]
            ids, logp = outs[i]
            args = _lib.DecodeArgs(group=G, n_sample=ns, start_token=START, end_token=END, pad_token=0, max_len=max_len,
                                   seed=seed, seq_id_base=lo * G, flags=_lib.DECODE_NO_EARLY_STOP)
            args.ids_out = ids.data_ptr()
            args.logp_out = logp.data_ptr() if ns else None
            _lib.check(e._h, lib.capdec_prefill(e._h, mem[lo:hi].data_ptr(), hi - lo, S, s.cuda_stream))
            _lib.check(e._h, lib.capdec_decode(e._h, C.byref(args), s.cuda_stream))
        for s in streams:
            ev = torch.cuda.Event(); ev.record(s); main.wait_event(ev)

    for i in range(3): run(i)
    ts = []
    for i in range(a.reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); run(10 + i); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    run(77); torch.cuda.synchronize()
    ids = torch.cat([o[0] for o in outs]).cpu()
    if ref_ids is None: ref_ids = ids
    same = float((ids == ref_ids).all(dim=1).float().mean())
    ms = sorted(ts)[len(ts) // 2]
    print(f"{a.workload} lanes={lanes} prio={a.prio}: rollout {ms:8.3f} ms  step {1e3 * ms / (max_len - 1):8.1f} us  "
          f"{n_img * G / ms * 1e3:9.0f} captions/s  rows identical to lanes=1: {same:.4f}", flush=True)
    for e in engs: e.close()
