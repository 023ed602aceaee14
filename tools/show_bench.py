import json,sys
def one(tag, d, kern=True):
    rr = d.get('rollout_roofline', {})
    print(f"{tag}: {d['config']['workload'][:40]} N={d['n_gpus']} {d['dtype']} value={d['value']:.0f} cap/s  step={d['decode_step_us']:.1f}us  e2e={d['e2e']['value']:.0f} ({d['e2e']['value']/d['value']:.3f})  rollout_frac={rr.get('frac',0):.3f} launches={d['gpu_launches']} clocks={d.get('clocks')}")
    if d.get('verify') is not None: print("   verify:", d['verify'])
    if kern:
        for k,v in d.get('kernels',{}).items():
            print(f"   {k:22s} n={v['launches']:4d} us={v['us_per_launch']:8.2f} share={v['share']:.3f} GB/s={v['GBps'] or 0:8.1f}")
for f in sys.argv[1:]:
    try: d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e: print(f, "ERR", e); continue
    one(f, d)
    for k, v in d.get('sub', {}).items():
        if 'config' in v: one("  sub " + k, v, kern=False)
        else: print("  sub", k, {a: v[a] for a in ('value', 'ms_per_step', 'decode_step_us', 'clocks') if a in v})
    if 'cpu_baseline' in d: print("  cpu_baseline:", d['cpu_baseline'])
