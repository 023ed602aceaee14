import json,sys
for f in sys.argv[1:]:
    try: d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e: print(f, "ERR", e); continue
    print(f"{f}: value={d['value']:.0f} cap/s  step={d['decode_step_us']:.1f}us  e2e={d['e2e']['value']:.0f}  rollout_frac={d['rollout_roofline']['frac']:.3f} launches={d['gpu_launches']}")
    for k,v in d.get('kernels',{}).items():
        print(f"   {k:22s} n={v['launches']:4d} us={v['us_per_launch']:8.2f} share={v['share']:.3f} GB/s={v['GBps'] or 0:8.1f}")
