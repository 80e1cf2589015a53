B="python bench.py --steps 30 --warmup 3 --no-cpu-baseline --mode replicas"
for nf in 1 2; do
$B --in-flight $nf > gpurun_out/x_inflight_$nf.json 2>> gpurun_out/x.err; echo rc=$?
python -c "
import json,sys
d=json.loads(open('gpurun_out/x_inflight_$nf.json').read().strip().splitlines()[-1])
print($nf, {k:d.get(k) for k in ('value','ms_per_step','e2e','latency_ms','proofs_in_flight')})
"
done
