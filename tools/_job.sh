python -m pytest tests -m gpu -x -q > gpurun_out/r1s13_pytest_gpu.log 2>&1; echo pytest=$?; tail -3 gpurun_out/r1s13_pytest_gpu.log
python bench.py --steps 20 --warmup 3 > gpurun_out/r1s13_bench.json 2> gpurun_out/r1s13_bench.err; echo bench=$?
for v in "g2_pair=2" "finalize_chunks=1" "g2_pair=2 finalize_chunks=1"; do
  o=""; for kv in $v; do o="$o --opt $kv"; done
  python bench.py --steps 20 --warmup 3 --no-cpu-baseline --mode replicas $o > "gpurun_out/r1s13_bench_$(echo $v | tr ' =' '__').json" 2>> gpurun_out/r1s13_bench.err; echo "variant $v rc=$?"
done
for f in gpurun_out/r1s13_bench*.json; do echo $f; python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1])
print({k:d.get(k) for k in ('value','ms_per_step','e2e','gpu_launches')}); print(d.get('roofline')); print(d.get('phases_ms') or d.get('kernels'))
"; done
