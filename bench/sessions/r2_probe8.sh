#!/bin/bash
# one 8-GPU box: scaling probe (world 8/4/2 x protocol, phase traces), multi-rank correctness, headline bench at 8
mkdir -p gpurun_out
tr() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) "$@"; }
tr bench/scale_probe.py wide_ann 200 8,4,2 tagged,barrier 2>gpurun_out/r2_probe8_err.log | grep '^{' > gpurun_out/r2_probe8_wide.jsonl
python - <<'PY'
import json
for l in open("gpurun_out/r2_probe8_wide.jsonl"):
    d = json.loads(l); print(d["config"], d["world"], d["protocol"], d["us_per_iteration"], "us/it", d["iterations_per_s"], "it/s", "E=%.12f" % d["last_err"])
PY
tr bench/scale_probe.py snn_mnist 2000 8,2 tagged 2>>gpurun_out/r2_probe8_err.log | grep '^{' > gpurun_out/r2_probe8_small.jsonl
tr bench/scale_probe.py xrd_snn 2000 8,2 tagged 2>>gpurun_out/r2_probe8_err.log | grep '^{' >> gpurun_out/r2_probe8_small.jsonl
python - <<'PY'
import json
for l in open("gpurun_out/r2_probe8_small.jsonl"):
    d = json.loads(l); print(d["config"], d["world"], d["protocol"], d["us_per_iteration"], "us/it", d["iterations_per_s"], "it/s")
PY
tr tests/mp_tp_check.py 2>&1 | grep -E "case|TP_CHECK|MISMATCH|Error|error" | tee gpurun_out/r2_tp_check_n8.log | tail -40
tr bench.py --gpus 8 --steps 8 --warmup 3 2>>gpurun_out/r2_probe8_err.log | grep '^{' > gpurun_out/r2_bench_n8.json
python -c 'import json; d=json.loads(open("gpurun_out/r2_bench_n8.json").read()); print("bench n8", d["value"], d["e2e"]["value"], d["detail"])'
tail -5 gpurun_out/r2_probe8_err.log
