#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_vranks.py tests/test_gpu_train.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -3
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 tests/mp_tp_check.py 2>gpurun_out/r2_call4_tp2.err | grep -E "TP_CHECK|MISMATCH|symmetric" | tail -5
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench/scale_probe.py wide_ann 200 2 barrier,tagged 2>gpurun_out/r2_call4_probe.err | grep '^{' > gpurun_out/r2_probe2b_wide.jsonl
python -c '
import json
for l in open("gpurun_out/r2_probe2b_wide.jsonl"):
    d=json.loads(l); print(d["world"], d["protocol"], d["us_per_iteration"], "us/it E=%.12f" % d["last_err"])
    for r,ph in enumerate(d["phase_cycles_by_rank"]): print("  rank",r,ph)'
tail -3 gpurun_out/r2_call4_probe.err
