#!/bin/bash
mkdir -p gpurun_out
tr() { timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) "$@"; }
for mb in 0 64 24; do
HPNN_PREFETCH_MB=$mb tr bench/scale_probe.py wide_ann 200 8 barrier 2>>gpurun_out/r2_probe8f_err.log | grep '^{' > gpurun_out/r2_probe8f_pf$mb.jsonl
python - gpurun_out/r2_probe8f_pf$mb.jsonl <<'PY'
import json,sys
for l in open(sys.argv[1]):
    d=json.loads(l); print(sys.argv[1].split("/")[-1], d["us_per_iteration"], "us/it", d["iterations_per_s"], "it/s E=%.12f" % d["last_err"]); print("   ", d["phase_cycles"])
PY
done
grep -v "^\*\|OMP_NUM\|^$" gpurun_out/r2_probe8f_err.log | tail -3 | cut -c1-300
