#!/bin/sh
# Stage the UNMODIFIED reference hot-path sources + Beauty data + checkpoint under baseline/_ref/ (git-ignored; it
# travels to the GPU box with gpurun) so tests/test_gpu_reference_trainer_dropin.py can run the reference's own
# Trainer / main.py loop with MODEL_DICT["bsarec"] swapped.  Build container only; remove it with
# `rm -rf baseline/_ref` when the evidence run is done — nothing in the product or in bench.py reads it.
set -e
SRC=${1:-/root/reference/src}
DST="$(dirname "$0")/_ref/src"
rm -rf "$DST"
mkdir -p "$DST/data" "$DST/output"
cp "$SRC"/main.py "$SRC"/trainers.py "$SRC"/utils.py "$SRC"/dataset.py "$SRC"/metrics.py "$DST"/
cp -r "$SRC"/model "$DST"/model
cp "$SRC"/data/Beauty.txt "$DST"/data/
cp "$SRC"/output/BSARec_Beauty_best.pt "$DST"/output/
echo "staged $(find "$DST" -type f | wc -l) files under $DST"
