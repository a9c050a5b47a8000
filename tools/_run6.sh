python tools/fan_check.py 2>&1 | tail -6
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s9_pytest_fan.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/s9_pytest_fan.log
for c in 2 1 3 5; do
  timeout 600 python bench.py --config $c --no-cpu-baseline --steps 20 > gpurun_out/s9d_bench_c$c.json 2> gpurun_out/s9d_bench_c$c.err; echo rc=$?
done
RT_PK_FAN=0 timeout 600 python bench.py --config 2 --no-cpu-baseline --steps 20 > gpurun_out/s9d_bench_c2_nofan.json 2>/dev/null
python -c "
import json,glob
for f in sorted(glob.glob('gpurun_out/s9d_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d.get('ms_per_frame'),4), round(d.get('value')), round(d['roofline']['frac'],3), round(d['e2e']['ms_per_step'],3))
    except Exception as e: print(f, 'ERR', e)
"
