#!/usr/bin/env python
"""profiles/r01_traffic.json from the ncu --set full summaries: dram bytes (read + write) per launch and kernel tag.
    python tools/traffic_table.py"""
import csv, json, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = {"_source": "profiles/r01_v2_ncu_full_cfg{2,3,5}.csv (ncu --set full --clock-control none, one decode "
                  "step; dram__bytes_read.sum + dram__bytes_write.sum per launch, median over the step's launches of that kernel)"}
for w in ("cfg2", "cfg3", "cfg5"):
    rows = list(csv.reader(open(os.path.join(ROOT, "profiles", f"r01_v2_ncu_full_{w}.csv"))))
    h = rows[0]
    ir = [i for i, x in enumerate(h) if x.startswith("dram__bytes_read")][0]
    iw = [i for i, x in enumerate(h) if x.startswith("dram__bytes_write")][0]
    unit = lambda s: 1e6 if "Mbyte" in s else (1e3 if "Kbyte" in s else 1)
    tags, prev = {}, None
    for r in rows[1:]:
        n = r[0]
        if "SwapQKV" in n: t = "gemm_qkv"
        elif "self_attn" in n: t = "self_attn"
        elif "cross_attn" in n: t = "cross_attn"
        elif "SwapLinear<bf16, 0>" in n: t = "gemm_cross_q"
        elif "SwapLinear<bf16, 1>" in n: t = "gemm_ffn1"
        elif "SwapLinear<float" in n: t = "gemm_ffn2_part"
        elif "sum_ln" in n: t = "sum_ln"
        elif "vocab_head" in n: t = "vocab_head"
        elif "SwapNoLN" in n:  # order inside a layer: self-out, cross-out, FFN2
            t = "gemm_self_out_ln" if prev == "self_attn" else ("gemm_cross_out_ln" if prev == "cross_attn" else "gemm_ffn2_ln")
        else: t = n
        prev = t
        tags.setdefault(t, []).append(float(r[ir]) * unit(h[ir]) + float(r[iw]) * unit(h[iw]))
    out[w] = {t: int(sorted(v)[len(v) // 2]) for t, v in tags.items()}
json.dump(out, open(os.path.join(ROOT, "profiles", "r01_traffic.json"), "w"), indent=1)
print(json.dumps(out, indent=1))
