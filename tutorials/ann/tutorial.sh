#!/bin/bash
# RRUFF powder-XRD tutorial (reference: tutorials/ann/tutorial.bash): 851 inputs (temperature +
# 850 bins of the diffraction pattern) -> 230 hidden -> 230 space groups, BPM.
#   1. unpack the RRUFF "XY_RAW" and "DIF" archives into ./rruff/raw and ./rruff/dif (no
#      download is attempted); without them a synthetic stand-in of the same shape is used;
#   2. pdif integrates every pattern into a sample file;
#   3. one training pass + ROUNDS restarts from kernel.opt, then run_nn on the training set
#      itself (as the reference tutorial does) and the PASS count is printed.
# usage: tutorial.sh [ROUNDS=2] [N_SYNTH=64] [type=ANN|SNN]
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"; ROOT="$(cd "$HERE/../.." && pwd)"
ROUNDS=${1:-2}; NS=${2:-64}; TYPE=${3:-ANN}
BIN="$ROOT/build/bin"
[ -x "$BIN/train_nn" ] || python "$ROOT/hpnn_b200/build.py" -q
mkdir -p samples
if [ -d rruff/dif ] && [ -d rruff/raw ]; then
  "$BIN/pdif" rruff -i 850 -o 230 -s samples
else
  echo "RRUFF files not found: generating a synthetic XRD-shaped stand-in"
  SNN=""; [ "$TYPE" = "SNN" ] && SNN="--snn"
  python "$ROOT/tools/synth_samples.py" --out samples --n "$NS" --inputs 851 --outputs 230 --clusters $SNN
fi
cat > nn.conf <<!
[name] RRUFF
[type] $TYPE
[init] generate
[seed] 10958
[input] 851
[hidden] 230
[output] 230
[train] BPM
[sample_dir] ./samples
[test_dir] ./samples
!
sed -e 's/^\[init\].*/[init] kernel.opt/' nn.conf > cont.conf
echo "training NN -- turn 0"; "$BIN/train_nn" -v -v nn.conf > log 2>&1
for r in $(seq 1 "$ROUNDS"); do echo "training NN -- turn $r"; "$BIN/train_nn" -v -v cont.conf > log 2>&1; done
"$BIN/run_nn" -v -v cont.conf > results 2>&1
echo "PASS: $(grep -c PASS results) / $(ls samples | wc -l)"
