for cap in 768 896 1152; do
  timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --prefill-cap $cap 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read())
f=lambda x:{k:(round(v,1) if isinstance(v,float) else v) for k,v in x.items() if k in ('value','ttft_ms_p50','ttft_ms_p99','itl_ms_p50','itl_ms_p99','wall_s')}
print('CAP',$cap,'ms/step',round(d['ms_per_step'],3),'e2e',f(d['e2e']),'eng',f(d['e2e_engine']),'ws',f(d.get('ws') or {}))
print('   slowest', d['e2e_engine'].get('slowest_steps_with_decode_rows'))"
done
