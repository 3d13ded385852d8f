#!/bin/bash
# Long optimisation run on MNIST with a softmax network and momentum (the role of the reference's
# tutorials/mnist/opt_mnist.bash): SNN 784-HIDDEN-10, BPM, one pass from a generated kernel and
# then ROUNDS-1 passes that restart from kernel.opt.  After every pass one line
#   ITER[i] PASS = x% OPT = y%      (test-set PASS rate, first-try OK rate of the training pass)
# is printed and appended to ./mnist/raw; while a pass runs, a progress bar is refreshed from the
# training log (the reference drives this with `watch`; here it is a background loop, so it also
# works without a terminal).  Expects the sample directories made by tutorial.sh (run that first,
# or let this script create the synthetic stand-in).
# usage: opt_mnist.sh [ROUNDS=30] [HIDDEN="1000 1000"] [N_SYNTH_TRAIN=200] [N_SYNTH_TEST=50]
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"; ROOT="$(cd "$HERE/../.." && pwd)"
ROUNDS=${1:-30}; HIDDEN=${2:-"1000 1000"}; NTR=${3:-200}; NTE=${4:-50}
BIN="$ROOT/build/bin"
[ -x "$BIN/train_nn" ] || python "$ROOT/hpnn_b200/build.py" -q
mkdir -p mnist/samples mnist/tests; cd mnist
if [ -z "$(ls samples 2>/dev/null | head -1)" ]; then
  if [ -f train_images ] && [ -f train_labels ] && [ -f test_images ] && [ -f test_labels ]; then
    "$BIN/pmnist" --scale ./samples ./tests
  else
    echo "no samples: generating a synthetic MNIST-shaped stand-in (SNN targets)"
    python "$ROOT/tools/synth_samples.py" --out ./samples --n "$NTR" --clusters --seed 1 --snn
    python "$ROOT/tools/synth_samples.py" --out ./tests --n "$NTE" --clusters --seed 2 --start 60001 --snn
  fi
fi
cat > mnist_snn.conf <<!
[name] MNIST
[type] SNN
[init] generate
[seed] 10958
[input] 784
[hidden] $HIDDEN
[output] 10
[train] BPM
[sample_dir] ./samples
[test_dir] ./tests
!
sed -e 's/^\[init\].*/[init] kernel.opt/' -e 's/^\[seed\].*/[seed] 1/' mnist_snn.conf > cont_mnist_snn.conf
NS=$(ls samples | wc -l); NT=$(ls tests | wc -l)
rate() { awk -v a="$1" -v b="$2" 'BEGIN{printf "%.1f", (b>0)?100*a/b:0}'; }
monitor() {  # progress of the running pass, one line every 5 s
  while sleep 5; do
    n=$(grep -c TRAINING log 2>/dev/null || true)
    awk -v n="$n" -v t="$NS" -v i="$1" 'BEGIN{f=(t>0)?n/t:0; b=""; for(k=0;k<10;k++) b=b ((k<int(10*f))?"#":"."); printf "ITER[%d] [%s](%.1f%%)\n", i, b, 100*f}'
  done
}
plot() {
  if command -v gnuplot >/dev/null; then
    gnuplot -e 'set term dumb size 80,30; set y2tics; set key below; plot "raw" u 1:2 w lp t "PASS" axis x1y1, "raw" u 1:3 w lp t "OPT" axis x1y2'
  else  # no gnuplot: a bar per pass
    awk '{n=int($2/2); b=""; for(k=0;k<n;k++) b=b "#"; printf "%3d PASS %5.1f%% |%s\n", $1, $2, b}' raw
  fi
}
: > raw
for i in $(seq 1 "$ROUNDS"); do
  conf=cont_mnist_snn.conf; [ "$i" -eq 1 ] && conf=mnist_snn.conf
  : > log
  monitor "$i" & MPID=$!
  "$BIN/train_nn" -v -v "$conf" > log 2>&1
  kill $MPID 2>/dev/null || true; wait $MPID 2>/dev/null || true
  "$BIN/run_nn" -v -v cont_mnist_snn.conf > results 2>&1
  XRS=$(rate "$(grep -c PASS results || true)" "$NT"); XOK=$(rate "$(grep -c ' OK' log || true)" "$NS")
  echo "$i $XRS $XOK" >> raw
  echo "ITER[$i] PASS = $XRS% OPT = $XOK%"
  [ "$i" -gt 1 ] && plot
done
echo "All DONE!"
