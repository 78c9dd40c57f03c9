#!/bin/bash
# First GPU call of the next session: the variants prepared (and not yet measured) at the end of round 1.
#   tools/gpu_next.sh build                      here (nvcc cross-compiles, ~1 min)
#   gpurun --timeout 120 -- 'tools/gpu_next.sh run'     on the B200 (about 15 s of run time; results in gpurun_out/ab_quick.txt)
# Each variant is first checked bit-equal against the default build on three renders (tools/ab_quick.py).
set -u
cd "$(dirname "$0")/.."
NAMES="base inap prerej prerej8 both8 coop coop8 coopprerej8"
if [ "${1:-}" = build ]; then
  VARIANTS="base: inap:-DPT_INSIDE_APPROX=1 prerej:-DPT_TRI_PREREJECT=1 prerej8:-DPT_TRI_PREREJECT=1,-DPT_TRAV_MIN_BLOCKS=8 \
both8:-DPT_TRI_PREREJECT=1,-DPT_INSIDE_APPROX=1,-DPT_TRAV_MIN_BLOCKS=8 coop:-DPT_COOP_EXACT=1 \
coop8:-DPT_COOP_EXACT=1,-DPT_TRAV_MIN_BLOCKS=8 coopprerej8:-DPT_COOP_EXACT=1,-DPT_TRI_PREREJECT=1,-DPT_TRAV_MIN_BLOCKS=8" \
    tools/ab_variants.sh build
else
  python tools/ab_quick.py $NAMES
fi
