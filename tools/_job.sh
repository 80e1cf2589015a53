python -m pytest tests -m gpu -x -q -k "msm" > gpurun_out/x_pytest_msm.log 2>&1; echo pytest=$?; tail -5 gpurun_out/x_pytest_msm.log
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --mode replicas"
$B > gpurun_out/x_default.json 2> gpurun_out/x.err; echo rc=$?
$B --opt reduce_fold=1 > gpurun_out/x_fold.json 2>> gpurun_out/x.err; echo rc=$?
for f in gpurun_out/x_default.json gpurun_out/x_fold.json; do echo $f; python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1])
print({k:d.get(k) for k in ('value','ms_per_step')}); print(d.get('phases_ms') or d.get('kernels'))
"; done
python tools/msm_sweep.py --min 16 --max 20 --g2-max 20 --reps 3 --opt reduce_fold=0 > gpurun_out/x_sweep_nofold.jsonl 2>>gpurun_out/x.err
python tools/msm_sweep.py --min 16 --max 20 --g2-max 20 --reps 3 --opt reduce_fold=1 > gpurun_out/x_sweep_fold.jsonl 2>>gpurun_out/x.err
python - <<'P'
import json
for f in ('gpurun_out/x_sweep_nofold.jsonl','gpurun_out/x_sweep_fold.jsonl'):
    for l in open(f):
        d=json.loads(l); print(f[-14:], d['group'], d['log_n'], round(d['ms'],3), d['in_exponent_ok'], d['phases_ms'].get('msm.reduce'), d['phases_ms'].get('msm.finalize'))
P
