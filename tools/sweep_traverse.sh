#!/bin/bash
# Tuning sweep of the traversal kernel's two constants (PT_NODE_BURST, PT_MIN_NODE_LANES) and of the samples per
# wavefront step.  Builds one library per variant here (nvcc cross-compiles), benchmarks them on the GPU box:
#   tools/sweep_traverse.sh build      (in the build container)
#   tools/sweep_traverse.sh run        (on the GPU box, via gpurun)
set -u
HERE=$(cd "$(dirname "$0")/.." && pwd)
V=$HERE/path-tracing_b200/lib/variants
VARIANTS="4:1 8:8 16:8 16:12 16:16 16:20 32:16 32:24"
if [ "${1:-}" = build ]; then
  mkdir -p $V
  for v in $VARIANTS; do
    b=${v%%:*}; m=${v##*:}
    nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -lineinfo -std=c++17 -Xcompiler -fPIC -shared \
      -DPT_NODE_BURST=$b -DPT_MIN_NODE_LANES=$m $HERE/path-tracing_b200/csrc/pt_api.cu -o $V/libptrender_b${b}_m${m}.so &
  done
  wait; ls -la $V
else
  mkdir -p $HERE/gpurun_out
  for v in $VARIANTS; do
    b=${v%%:*}; m=${v##*:}
    for q in 2 4; do
      PTB200_RENDER_SO=$V/libptrender_b${b}_m${m}.so python $HERE/bench.py --workload torus1m --spp 32 --steps 2 --warmup 1 --no-cpu --q-shift $q \
        2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('b$b m$m q$q', round(d['value']), 'Mrays/s', {k: round(v,2) for k,v in r['class_ms'].items()}, 'nodes/ray', round(r['nodes_per_ray'],2))"
    done
  done | tee $HERE/gpurun_out/sweep_traverse.txt
fi
