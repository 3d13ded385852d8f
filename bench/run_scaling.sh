#!/bin/bash
# usage: bench/run_scaling.sh N [steps] -- ours + reference arms of bench.py at N GPUs for the
# wide_ann and snn_mnist configs; JSON lines are appended to gpurun_out/scale_N.jsonl
N=${1:-2}; STEPS=${2:-4}
mkdir -p gpurun_out; OUT=gpurun_out/scale_$N.jsonl; : > $OUT
run() {  # impl config
  if [ "$N" -gt 1 ]; then
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port $((29600 + RANDOM % 300)) bench.py --gpus $N --impl $1 --config $2 --steps $STEPS --warmup 3 2>gpurun_out/scale_err.log | grep '^{' >> $OUT
  else
    timeout 900 python bench.py --gpus 1 --impl $1 --config $2 --steps $STEPS --warmup 3 2>gpurun_out/scale_err.log | grep '^{' >> $OUT
  fi
  echo "[$1 $2 N=$N] rc=$? $(tail -1 $OUT | python -c 'import sys,json
try:
  d=json.loads(sys.stdin.read()); print(round(d["value"],3), d["unit"], "iters/s", round(d["config"]["iterations_per_s"],1), "e2e", round(d["e2e"]["value"],3))
except Exception as e: print("no json", e)')"
}
if [ "$N" -gt 1 ]; then
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 tests/mp_tp_check.py 2>&1 | grep -E "case|TP_CHECK" 
fi
for cfg in wide_ann snn_mnist; do run ours $cfg; run reference $cfg; done
