#!/bin/bash
# usage: gpu_ll_ab.sh N  -- tagged-word exchange (exchange=peer) against the flag protocol (peer_flag) and NCCL on N GPUs
N=$1; D=gpurun_out/r2ll_n$N; mkdir -p $D
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 "$@"; }
run tests/mgpu_worker.py > $D/mgpu_worker.log 2>&1; echo "worker rc=$?"; tail -2 $D/mgpu_worker.log
for rep in 1 2; do for ex in peer peer_flag; do
  run bench.py --gpus $N --steps 20 --warmup 3 --no-cpu --no-e2e --exchange $ex > $D/c3_s20_${ex}_$rep.json 2> $D/err.log; echo "c3 $ex $rep rc=$?"
done; done
run bench.py --gpus $N --no-cpu --no-e2e --exchange peer > $D/c3_s100_peer.json 2>> $D/err.log
run bench.py --gpus $N --no-cpu --no-e2e --exchange peer_flag > $D/c3_s100_peer_flag.json 2>> $D/err.log
run bench.py --gpus $N --config 2 --no-cpu --no-e2e --exchange peer > $D/c2_peer.json 2>> $D/err.log
run bench.py --gpus $N --config 2 --no-cpu --no-e2e --exchange peer_flag > $D/c2_peer_flag.json 2>> $D/err.log
python - <<PY
import json,glob
for f in sorted(glob.glob('$D/c*.json')):
    try:
        j=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(j['value'],2), 'us/step', round(1e3*j['ms_per_step'],2), 'sweep_us', round(1e3*j['roofline']['launch_ms'],2), 'err', j['config'].get('last_err'))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 $D/err.log
