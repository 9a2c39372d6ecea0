for m in 0 1 2 4 7; do
LEGO_DBG_SKIP=$m python bench.py --stages 3 --lanes 296 --handles 1 --steps 12 --warmup 4 --no-e2e --no-cpu-baseline --profile-json gpurun_out/skip_$m.json > gpurun_out/skip_$m.log 2>&1
done
python - <<'PY'
import json
for m in (0,1,2,4,7):
    try:
        d=json.load(open('gpurun_out/skip_%d.json'%m))
        ks=d['per_kernel_ms_over_timed_steps']
        print(m, [(k, round(v['ms'],2) if isinstance(v,dict) and 'ms' in v else v) for k,v in (ks.items() if isinstance(ks,dict) else []) if 'extract' in k])
    except Exception as e: print(m, 'err', e)
PY
