for v in "" _nofan _nopf _nofanpf; do
  for c in 1 2; do
    RT_B200_LIB=$PWD/opengl_raytracing_b200/csrc/librt_b200$v.so timeout 300 python bench.py --config $c --no-cpu-baseline --steps 30 > gpurun_out/s9e_bench_c${c}_v$v.json 2>/dev/null; echo rc=$?
  done
done
python -c "
import json,glob
for f in sorted(glob.glob('gpurun_out/s9e_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d.get('ms_per_frame'),4), round(d.get('value')))
    except Exception as e: print(f, 'ERR', e)
"
