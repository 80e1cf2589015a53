python -m pytest tests -m gpu -x -q > gpurun_out/x_pytest.log 2>&1; echo pytest=$?; tail -3 gpurun_out/x_pytest.log
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --mode replicas"
$B > gpurun_out/x_default.json 2> gpurun_out/x.err; echo rc=$?
for f in gpurun_out/x_default.json; do echo $f; python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1])
print({k:d.get(k) for k in ('value','ms_per_step')}); print(d.get('phases_ms') or d.get('kernels'))
"; done
python tools/msm_sweep.py --min 16 --max 20 --groups 1 --reps 3 > gpurun_out/x_sweep_g1.jsonl 2>>gpurun_out/x.err
python tools/msm_sweep.py --min 16 --max 17 --groups 1 --reps 3 --chunk 16 --no-check > gpurun_out/x_sweep_g1_c16.jsonl 2>>gpurun_out/x.err
python tools/msm_sweep.py --min 16 --max 17 --groups 1 --reps 3 --chunk 32 --no-check > gpurun_out/x_sweep_g1_c32.jsonl 2>>gpurun_out/x.err
for ch in 0 32 48 64 80; do
python tools/msm_sweep.py --min 18 --max 20 --g2-max 20 --groups 2 --reps 3 --chunk $ch --no-check > gpurun_out/x_sweep_g2_c$ch.jsonl 2>>gpurun_out/x.err
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/x_sweep_g*.jsonl')):
    for l in open(f):
        d=json.loads(l); print(f[17:], d['group'], d['log_n'], round(d['ms'],3), d['in_exponent_ok'], {k[4:]:v for k,v in d['phases_ms'].items()})
P
