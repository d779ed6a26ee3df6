#!/usr/bin/env bash
# TEST / BASELINE INFRASTRUCTURE (see oracle/__init__.py) -- not product code.
#
# Stages the UNMODIFIED reference modules the CPU arm of bench.py times into oracle/_ref/ (git-ignored, so the
# reference's sources never enter this repository's history; NOT gpurun-ignored, so the staged files travel to the
# GPU box like the built .so):
#     phase1_lifting/baselineModel.py   LinearModel / weight_init        (the module under test)
#     phase5_loop/losses.py             loss_MPJPE / TriangleLoss*       (the losses around it)
# and records their sha256 in oracle/_ref/SHA256 so that bench.py can state which bytes it ran.  The reference is pure
# Python: there is nothing to compile; `bench.py --impl reference` and `cpu_baseline` import these files as they are
# (kind = "reference").  Re-run after the reference changes:   bash oracle/make_ref.sh [/root/reference]
set -euo pipefail
REF="${1:-${LIFTER_REFERENCE:-/root/reference}}"
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
if [ ! -d "$REF/phase1_lifting" ]; then
    echo "make_ref: $REF/phase1_lifting not found (the reference only exists in the build container)" >&2
    exit 3
fi
mkdir -p "$OUT"
install -m 0644 "$REF/phase1_lifting/baselineModel.py" "$OUT/baselineModel.py"
install -m 0644 "$REF/phase5_loop/losses.py" "$OUT/losses.py"
( cd "$OUT" && sha256sum baselineModel.py losses.py > SHA256 )
cat "$OUT/SHA256"
