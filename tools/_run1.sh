set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/s9_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s9_pytest_gpu.log
tail -3 gpurun_out/s9_pytest_gpu.log
timeout 300 python bench.py > gpurun_out/s9_bench_c2.json 2> gpurun_out/s9_bench_c2.err; echo rc=$?
for v in _w16 _w4 _w4c5 ""; do
  for c in 2 1; do
    RT_B200_LIB=$PWD/opengl_raytracing_b200/csrc/librt_b200$v.so timeout 300 python bench.py --config $c --no-cpu-baseline --steps 30 > gpurun_out/s9_bench_c${c}_v$v.json 2> gpurun_out/s9_bench_c${c}_v$v.err; echo rc=$?
  done
done
python -c "
import json,glob
for f in sorted(glob.glob('gpurun_out/s9_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d.get('ms_per_frame'), d.get('value'), d['e2e']['ms_per_step'])
    except Exception as e: print(f, 'ERR', e)
"
