#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_dp.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -8
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 tests/mp_tp_check.py 2>gpurun_out/r2_call11_tp2.err | grep -E "dp-merge|TP_CHECK|MISMATCH|symmetric" | tail -5
tail -4 gpurun_out/r2_call11_tp2.err | cut -c1-300
for cfg in snn_mnist wide_ann; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29614 bench/dp_merge.py $cfg 10 2>>gpurun_out/r2_call11_dp.err | grep '^{' | tee -a gpurun_out/r2_dp_merge_n2.jsonl | cut -c1-330
done
timeout 200 python bench/dp_merge.py wide_ann 10 factored 2>>gpurun_out/r2_call11_dp.err | grep '^{' | tee gpurun_out/r2_dp_merge_n1.jsonl | cut -c1-330
grep -v "^\*\|OMP_NUM\|^$" gpurun_out/r2_call11_dp.err | tail -5 | cut -c1-300
