#!/usr/bin/env python
"""profiles/ncu_kernels.json from the ncu --set full summaries (profiles/r02_ncu_full_cfg{2,3}.csv): per kernel tag the
median gpu__time_duration (us) and dram bytes (read + write) per launch.  bench.py reports them beside its own
event-pair timings (`ncu_us_per_launch`, `roofline.traffic`).
    python tools/traffic_table.py"""
import csv, json, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = {"_source": "profiles/r02_ncu_full_cfg2.csv, r02_ncu_full_cfg3_final.csv: ncu --set full --clock-control none on `bench.py --workload W --steps 1 "
                  "--warmup 3 --no-cpu --no-sub`, launches of the LAST decode step of a rollout (cached length 48-49: the "
                  "self-attention figure is the largest of the rollout); median over the captured launches of each kernel"}
unit = lambda s: 1e6 if "Mbyte" in s else (1e3 if "Kbyte" in s else 1)
for w in ("cfg2", "cfg3"):
    final = os.path.join(ROOT, "profiles", f"r02_ncu_full_{w}_final.csv")  # (cfg3: re-captured after the last kernel changes)
    rows = list(csv.reader(open(final if os.path.exists(final) else os.path.join(ROOT, "profiles", f"r02_ncu_full_{w}.csv"))))
    h = rows[0]
    it = [i for i, x in enumerate(h) if x.startswith("gpu__time_duration")][0]
    ir = [i for i, x in enumerate(h) if x.startswith("dram__bytes_read")][0]
    iw = [i for i, x in enumerate(h) if x.startswith("dram__bytes_write")][0]
    tags, prev = {}, None
    for r in rows[1:]:
        n = r[0]
        if "WsQKV" in n or "SwapQKV" in n: t = "gemm_qkv"
        elif "self_attn" in n: t = "self_attn"
        elif "cross_attn" in n: t = "cross_attn"
        elif "WsLinear<0>" in n or "SwapLinear<bf16, 0>" in n: t = "gemm_cross_q"
        elif "WsLinear<1>" in n or "SwapLinear<bf16, 1>" in n: t = "gemm_ffn1"
        elif "SwapLinear<float" in n: t = "gemm_ffn2_part"
        elif "sum_ln" in n: t = "sum_ln"
        elif "vocab_head" in n: t = "vocab_head"
        elif "gemm_umma_kernel" in n and "float" in n: t = "gemm_vocab"
        elif n.startswith("head_kernel"): t = "head"
        elif "chain_kernel" in n: t = {"0": "chain_in_proj", "1": "chain_attn_out", "2": "chain_ffn"}[n.split(",")[-1].strip(" >")]
        elif "SwapNoLN" in n:  # order inside a layer: self-out, cross-out, FFN2
            t = "gemm_self_out_ln" if prev == "self_attn" else ("gemm_cross_out_ln" if prev == "cross_attn" else "gemm_ffn2_ln")
        else: t = n
        prev = t
        tags.setdefault(t, []).append((float(r[it]), float(r[ir]) * unit(h[ir]) + float(r[iw]) * unit(h[iw])))
    med = lambda v: sorted(v)[len(v) // 2]
    out[w] = {t: {"us": med([a for a, _ in v]), "dram_bytes": int(med([b for _, b in v])), "launches": len(v)} for t, v in tags.items()}
out["cfg4"] = out["cfg3"]  # (the sharded configuration runs the same kernels on fewer rows; figures are the unsharded ones)
json.dump(out, open(os.path.join(ROOT, "profiles", "ncu_kernels.json"), "w"), indent=1)
print(json.dumps(out, indent=1))
