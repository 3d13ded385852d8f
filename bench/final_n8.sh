#!/bin/bash
# one 8-GPU box: correctness check, then the two benchmark configs at 8 GPUs (JSON lines -> gpurun_out/final_n8.jsonl)
mkdir -p gpurun_out; OUT=gpurun_out/final_n8.jsonl; : > $OUT
tr() { timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) "$@"; }
tr tests/mp_tp_check.py 2>&1 | grep -E "case|TP_CHECK"
summ() { tail -1 $OUT | python -c 'import sys,json
d=json.loads(sys.stdin.read()); print(sys.argv[1], round(d["value"],3), d["unit"], "it/s", round(d["config"]["iterations_per_s"],1), "e2e", round(d["e2e"]["value"],3))' "$1"; }
tr bench.py --gpus 8 --config wide_ann --steps 8 --warmup 3 2>gpurun_out/n8_err.log | grep '^{' >> $OUT; summ wide_ann
HPNN_PREFETCH_MB=48 tr bench.py --gpus 8 --config wide_ann --steps 8 --warmup 3 2>>gpurun_out/n8_err.log | grep '^{' >> $OUT; summ wide_ann_prefetch48
HPNN_LAZY=0 tr bench.py --gpus 8 --config wide_ann --steps 8 --warmup 3 2>>gpurun_out/n8_err.log | grep '^{' >> $OUT; summ wide_ann_nolazy
tr bench.py --gpus 8 --config snn_mnist --steps 16 --warmup 3 2>>gpurun_out/n8_err.log | grep '^{' >> $OUT; summ snn_mnist
timeout 300 python -m pytest tests/test_gpu_capi.py -x -q 2>&1 | tail -2
