#!/bin/bash
# MNIST tutorial (reference: tutorials/mnist/tutorial.bash): ANN 784-300-10, BP.
#   1. put the four MNIST idx files into ./mnist as train_images train_labels test_images
#      test_labels (no download is attempted: the build boxes have no network); without them
#      a synthetic clustered stand-in of the same shape is generated;
#   2. pmnist converts them to sample files;
#   3. one training pass with a generated kernel, then ROUNDS passes that restart from
#      kernel.opt; after every pass the PASS rate on the test set and the first-try OK rate
#      are appended to ./mnist/raw (columns: round, PASS %, OK %).
# usage: tutorial.sh [ROUNDS=3] [N_SYNTH_TRAIN=200] [N_SYNTH_TEST=50]
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"; ROOT="$(cd "$HERE/../.." && pwd)"
ROUNDS=${1:-3}; NTR=${2:-200}; NTE=${3:-50}
BIN="$ROOT/build/bin"
[ -x "$BIN/train_nn" ] || python "$ROOT/hpnn_b200/build.py" -q
mkdir -p mnist/samples mnist/tests; cd mnist
if [ -f train_images ] && [ -f train_labels ] && [ -f test_images ] && [ -f test_labels ]; then
  "$BIN/pmnist" --scale ./samples ./tests
else
  echo "MNIST files not found: generating a synthetic MNIST-shaped stand-in"
  python "$ROOT/tools/synth_samples.py" --out ./samples --n "$NTR" --clusters --seed 1
  python "$ROOT/tools/synth_samples.py" --out ./tests --n "$NTE" --clusters --seed 2 --start 60001
fi
cat > mnist_ann.conf <<!
[name] MNIST
[type] ANN
[init] generate
[seed] 10958
[input] 784
[hidden] 300
[output] 10
[train] BP
[sample_dir] ./samples
[test_dir] ./tests
!
sed -e 's/^\[init\].*/[init] kernel.opt/' -e 's/^\[seed\].*/[seed] 1/' mnist_ann.conf > cont_mnist_ann.conf
NS=$(ls samples | wc -l); NT=$(ls tests | wc -l)
rate() { awk -v a="$1" -v b="$2" 'BEGIN{printf "%.1f", (b>0)?100*a/b:0}'; }
: > raw
"$BIN/train_nn" -v -v mnist_ann.conf > log 2>&1
"$BIN/run_nn" -v -v cont_mnist_ann.conf > results 2>&1
echo "0 $(rate $(grep -c PASS results) $NT) $(rate $(grep -c ' OK' log) $NS)" >> raw
for r in $(seq 1 "$ROUNDS"); do
  "$BIN/train_nn" -v -v cont_mnist_ann.conf > log 2>&1
  "$BIN/run_nn" -v -v cont_mnist_ann.conf > results 2>&1
  echo "$r $(rate $(grep -c PASS results) $NT) $(rate $(grep -c ' OK' log) $NS)" >> raw
done
echo "round PASS% OK%"; cat raw
command -v gnuplot >/dev/null && gnuplot -e 'set term dumb size 80,30; plot "raw" u 1:2 w lp t "PASS", "raw" u 1:3 w lp t "OPT"' || true
echo "All DONE!"
