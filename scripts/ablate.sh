#!/bin/bash
# Ablations / A-B runs of the tiled kernel on the EXPERIMENTS build (scripts/libpva_preproc_exp.so,
# `make -C pytorchvideo-accelerate_b200/csrc experiments`).  Each argument is one environment setting
# ("PVA_TILED_DEBUG=4", "PVA_TILED_MAXGRID=74 PVA_TILED_PDL=0", ...); prints us per launch for K=20 and
# K=200 back-to-back launches (scripts/steps_fit.py).  PVA_TILED_DEBUG bits: 1 no stores, 2 no loads,
# 4 no arithmetic.  Example:
#   scripts/ablate.sh "X=0" "PVA_TILED_DEBUG=4" "PVA_TILED_DEBUG=6" "PVA_TILED_DEBUG=5" "PVA_TILED_HALF_GRID=0"
export PVA_LIBRARY=scripts/libpva_preproc_exp.so
for v in "$@"; do
  echo "== $v"
  env $v python scripts/steps_fit.py --ks 20,200 --trials 8 --tag _abl 2>&1 | grep '"K"' | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['K'], round(d['per_step_us_median'], 2))"
done
