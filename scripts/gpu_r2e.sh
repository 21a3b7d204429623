#!/bin/bash
# 8 GPUs: multi-rank parity worker + bench lines of every config at N = 8 (config 3 also with the driver's 20-step window)
mkdir -p gpurun_out/r2e
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 "$@"; }
run tests/mgpu_worker.py > gpurun_out/r2e/mgpu8_worker.log 2>&1; echo "worker rc=$?"; tail -3 gpurun_out/r2e/mgpu8_worker.log
run bench.py --gpus 8 --steps 20 --warmup 3 --no-cpu > gpurun_out/r2e/bench_c3_n8_steps20.json 2> gpurun_out/r2e/bench_c3_n8_steps20.err; echo "c3 s20 rc=$?"; tail -c 250 gpurun_out/r2e/bench_c3_n8_steps20.json
run bench.py --gpus 8 --no-cpu > gpurun_out/r2e/bench_c3_n8.json 2> gpurun_out/r2e/bench_c3_n8.err; echo "c3 rc=$?"
run bench.py --gpus 8 --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2e/bench_c3_n8_steps20_b.json 2> /dev/null; echo "c3 s20 b rc=$?"
for c in 4 5 2; do
  run bench.py --gpus 8 --config $c --no-cpu > gpurun_out/r2e/bench_c${c}_n8.json 2> gpurun_out/r2e/bench_c${c}_n8.err; echo "c$c rc=$?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2e/bench_*.json')):
    try:
        j=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(j['value'],2), 'ms/step', round(j['ms_per_step'],4), 'frac', round(j['roofline']['frac'],4))
    except Exception as e: print(f, 'ERR', e)
PY
