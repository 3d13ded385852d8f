#!/bin/bash
mkdir -p gpurun_out
timeout 180 python -m pytest tests/test_gpu_umma.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -12
timeout 120 python -m pytest tests/test_run_kernel.py tests/test_gpu_dp.py -m gpu -q -p no:cacheprovider 2>&1 | tail -5
timeout 300 python bench/infer_sweep.py 2>&1 | tail -8 | cut -c1-400
cp gpurun_out/infer_sweep.json gpurun_out/r2_infer_sweep.json
timeout 300 python bench/ours_run_nn.py 2048 2>&1 | tail -5 | tee gpurun_out/r2_run_nn_e2e.jsonl
