python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 tools/msm_sweep.py --from 20 --to 26 --g2-max 24 --reps 3 > gpurun_out/r1_msm_sweep_b200_n8.raw 2> gpurun_out/x.err; echo sweep=$?
grep '^{' gpurun_out/r1_msm_sweep_b200_n8.raw > gpurun_out/r1_msm_sweep_b200_n8.jsonl
python - <<'P'
import json
for l in open('gpurun_out/r1_msm_sweep_b200_n8.jsonl'):
    d=json.loads(l); print(d['group'], d['log_n'], d['n_gpus'], round(d['ms'],3), round(d['mpoints_per_s'],1), round(d['frac_of_imad_peak'],3), d['in_exponent_ok'], d['phases_ms'].get('msm.exchange'))
P
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r1_bench_prove_2p20_n8.json 2>> gpurun_out/x.err; echo bench8=$?
python -c "
import json
d=json.loads([l for l in open('gpurun_out/r1_bench_prove_2p20_n8.json') if l.startswith('{')][-1])
print({k:d.get(k) for k in ('value','ms_per_step','n_gpus','e2e','sharded')})
"
tail -3 gpurun_out/x.err
