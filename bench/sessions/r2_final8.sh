#!/bin/bash
# one 8-GPU box: correctness log at 8 ranks, data-parallel merge, row-split batched inference, small-net fallback
mkdir -p gpurun_out
tr() { timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) "$@"; }
tr tests/mp_tp_check.py 2>gpurun_out/r2_final8_err.log | grep -E "case|TP_CHECK|MISMATCH" | tee gpurun_out/r2_tp_check_n8_final.log | tail -30
tr bench/dp_merge.py wide_ann 10 factored,nvls,nccl 2>>gpurun_out/r2_final8_err.log | grep '^{' | tee gpurun_out/r2_dp_merge_n8.jsonl | cut -c1-300
tr bench/infer_sweep_tp.py 2>>gpurun_out/r2_final8_err.log | grep '^{' | tee gpurun_out/r2_infer_tp_n8_mc.jsonl | cut -c1-260
HPNN_TP_BATCH_MC=0 tr bench/infer_sweep_tp.py 2>>gpurun_out/r2_final8_err.log | grep '^{' | tee gpurun_out/r2_infer_tp_n8_nomc.jsonl | tail -3 | cut -c1-260
tr bench.py --gpus 8 --config snn_mnist --steps 16 --warmup 3 2>>gpurun_out/r2_final8_err.log | grep '^{' > gpurun_out/r2_bench_snn_n8.json
python -c 'import json; d=json.loads(open("gpurun_out/r2_bench_snn_n8.json").read()); print("snn_mnist n8", d["value"], d["e2e"]["value"], d["detail"]["iterations_per_s"])'
grep -v "^\*\|OMP_NUM\|^$" gpurun_out/r2_final8_err.log | tail -4 | cut -c1-300
