#!/bin/bash
# usage: bench/final_n.sh N -- the two benchmark configs at N GPUs (JSON lines -> gpurun_out/final_nN.jsonl)
N=$1; mkdir -p gpurun_out; OUT=gpurun_out/final_n$N.jsonl; : > $OUT
tr() { timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) "$@"; }
summ() { tail -1 $OUT | python -c 'import sys,json
d=json.loads(sys.stdin.read()); print(sys.argv[1], round(d["value"],3), d["unit"], "it/s", round(d["config"]["iterations_per_s"],1), "e2e", round(d["e2e"]["value"],3))' "$1"; }
tr bench.py --gpus $N --config wide_ann --steps 8 --warmup 3 2>gpurun_out/n${N}_err.log | grep '^{' >> $OUT; summ wide_ann
tr bench.py --gpus $N --config snn_mnist --steps 16 --warmup 3 2>>gpurun_out/n${N}_err.log | grep '^{' >> $OUT; summ snn_mnist
