#!/bin/bash
mkdir -p gpurun_out
tr() { timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) "$@"; }
tr bench/scale_probe.py wide_ann 200 8 barrier 2>gpurun_out/r2_probe8c_err.log | grep '^{' > gpurun_out/r2_probe8c_default.jsonl
python - <<'PY'
import json
for l in open("gpurun_out/r2_probe8c_default.jsonl"):
    d=json.loads(l); print(d["world"], d["protocol"], d["us_per_iteration"], "us/it", d["iterations_per_s"], "it/s E=%.12f" % d["last_err"]); print("   ", d["phase_cycles"])
PY
tr bench.py --gpus 8 --steps 8 --warmup 3 2>>gpurun_out/r2_probe8c_err.log | grep '^{' > gpurun_out/r2_bench_n8c.json
python -c 'import json; d=json.loads(open("gpurun_out/r2_bench_n8c.json").read()); print("bench n8", d["value"], d["e2e"]["value"], d["detail"])'
grep -v "^\*\|OMP_NUM\|^$" gpurun_out/r2_probe8c_err.log | tail -5
