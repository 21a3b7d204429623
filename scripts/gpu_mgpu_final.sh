#!/bin/bash
# usage: gpu_mgpu_final.sh N   -- N GPUs: multi-rank parity worker + bench lines of every config (config 3 also with the driver's 20-step window)
N=$1; D=gpurun_out/r2f_n$N; mkdir -p $D
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 "$@"; }
run tests/mgpu_worker.py > $D/mgpu_worker.log 2>&1; echo "worker rc=$?"; tail -2 $D/mgpu_worker.log
run bench.py --gpus $N --steps 20 --warmup 3 --no-cpu > $D/bench_c3_steps20.json 2> $D/bench_c3_steps20.err; echo "c3 s20 rc=$?"
run bench.py --gpus $N --no-cpu --no-e2e > $D/bench_c3.json 2> $D/bench_c3.err; echo "c3 rc=$?"
run bench.py --gpus $N --steps 20 --warmup 3 --no-cpu --no-e2e > $D/bench_c3_steps20_b.json 2> /dev/null; echo "c3 s20 b rc=$?"
run bench.py --gpus $N --steps 20 --warmup 3 --no-cpu --no-e2e --exchange nccl > $D/bench_c3_steps20_nccl.json 2> /dev/null; echo "c3 s20 nccl rc=$?"
for c in 4 5 2; do
  run bench.py --gpus $N --config $c --no-cpu > $D/bench_c${c}.json 2> $D/bench_c${c}.err; echo "c$c rc=$?"
done
python - <<PY
import json,glob
for f in sorted(glob.glob('$D/bench_*.json')):
    try:
        j=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(j['value'],2), 'ms/step', round(j['ms_per_step'],4), 'frac', round(j['roofline']['frac'],4), 'sweep_ms', round(j['roofline']['launch_ms'],4))
    except Exception as e: print(f, 'ERR', e)
PY
