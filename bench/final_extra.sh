#!/bin/bash
# usage: bench/final_extra.sh N -- BASELINE configs 3 and 5 at N GPUs: xrd_snn training, run_nn batch sweep
N=$1; mkdir -p gpurun_out
tr() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) "$@"; }
tr bench.py --gpus $N --config xrd_snn --steps 16 --warmup 3 2>gpurun_out/extra_n${N}_err.log | grep '^{' > gpurun_out/xrd_n$N.json
python -c 'import sys,json
d=json.loads(open(sys.argv[1]).read()); print("xrd_snn", d["n_gpus"], round(d["value"],2), "samples/s", round(d["config"]["iterations_per_s"],1), "it/s")' gpurun_out/xrd_n$N.json
tr bench/infer_sweep_tp.py 2>>gpurun_out/extra_n${N}_err.log | grep '^{' | tee gpurun_out/infer_tp_n$N.jsonl
