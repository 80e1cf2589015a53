python -m pytest tests -m gpu -x -q > gpurun_out/r1_pytest_gpu_final.log 2>&1; echo pytest=$?; tail -3 gpurun_out/r1_pytest_gpu_final.log
python __graft_entry__.py smoke > gpurun_out/r1_smoke.log 2>&1; echo smoke=$?; tail -2 gpurun_out/r1_smoke.log
python bench.py --steps 20 --warmup 3 > gpurun_out/r1_bench_prove_2p20.json 2> gpurun_out/x.err; echo bench=$?
python -c "
import json
d=json.loads([l for l in open('gpurun_out/r1_bench_prove_2p20.json') if l.startswith('{')][-1])
print({k:d.get(k) for k in ('metric','value','unit','ms_per_step','e2e','gpu_launches','cpu_baseline','clocks','vs_baseline','dtype','config')}); print(d['roofline']); print(d['phases_ms'])
"
python bench.py --log-m 24 --steps 5 --warmup 3 --no-cpu-baseline --mode replicas > gpurun_out/r1_bench_prove_2p24.json 2>> gpurun_out/x.err; echo rc24=$?
python -c "
import json
d=json.loads([l for l in open('gpurun_out/r1_bench_prove_2p24.json') if l.startswith('{')][-1])
print({k:d.get(k) for k in ('value','ms_per_step','e2e')}); print(d.get('phases_ms'))
"
python tools/timeline.py > gpurun_out/r1_timeline_prove_2p20.txt 2>>gpurun_out/x.err; head -40 gpurun_out/r1_timeline_prove_2p20.txt
tail -3 gpurun_out/x.err
