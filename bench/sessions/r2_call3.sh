#!/bin/bash
# 2-GPU box: full GPU test-suite (1 GPU part + the 2-GPU multi test), TP check at 2 (VMM / multicast path),
# barrier-vs-tagged probe at 2, reference multi-GPU diagnostic
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --maxfail=10 -p no:cacheprovider > gpurun_out/r2_call3_pytest.log 2>&1; tail -15 gpurun_out/r2_call3_pytest.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 tests/mp_tp_check.py 2>gpurun_out/r2_call3_tp2.err | grep -E "case|TP_CHECK|MISMATCH" | tee gpurun_out/r2_tp_check_n2.log | tail -40
tail -5 gpurun_out/r2_call3_tp2.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench/scale_probe.py wide_ann 200 2 barrier,tagged 2>gpurun_out/r2_call3_probe.err | grep '^{' > gpurun_out/r2_probe2_wide.jsonl
python -c '
import json
for l in open("gpurun_out/r2_probe2_wide.jsonl"):
    d=json.loads(l); print(d["world"], d["protocol"], d["us_per_iteration"], "us/it E=%.12f" % d["last_err"])'
HPNN_MULTICAST=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29613 bench/scale_probe.py wide_ann 200 2 barrier 2>>gpurun_out/r2_call3_probe.err | grep '^{' > gpurun_out/r2_probe2_wide_nomc.jsonl
python -c '
import json
for l in open("gpurun_out/r2_probe2_wide_nomc.jsonl"):
    d=json.loads(l); print("no-multicast", d["world"], d["protocol"], d["us_per_iteration"], "us/it")'
timeout 500 python baseline/diag_ref_multigpu.py 2 > gpurun_out/r2_ref_diag_n2.log 2>&1; head -60 gpurun_out/r2_ref_diag_n2.log
