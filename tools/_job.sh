python -m pytest tests -m gpu -x -q > gpurun_out/x_pytest.log 2>&1; echo pytest=$?; tail -15 gpurun_out/x_pytest.log | cut -c1-300
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --mode replicas"
$B > gpurun_out/x_default.json 2> gpurun_out/x.err; echo rc=$?
$B --opt shared_sort=0 > gpurun_out/x_noshare.json 2>> gpurun_out/x.err; echo rc=$?
for f in gpurun_out/x_default.json gpurun_out/x_noshare.json; do echo $f; python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1])
print({k:d.get(k) for k in ('value','ms_per_step')}); print(d.get('phases_ms') or d.get('kernels'))
"; done
tail -5 gpurun_out/x.err
