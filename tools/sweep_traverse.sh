#!/bin/bash
# Tuning sweep of the traversal kernel's compile-time constants.  Builds one library per variant here (nvcc
# cross-compiles), benchmarks them on the GPU box:
#   tools/sweep_traverse.sh build      (in the build container)
#   tools/sweep_traverse.sh run        (on the GPU box, via gpurun)
set -u
HERE=$(cd "$(dirname "$0")/.." && pwd)
V=$HERE/path-tracing_b200/lib/variants
VARIANTS="${VARIANTS:-8:8:10 8:8:12 8:8:14 8:8:16 8:8:8}"   # burst:min_lanes:min_blocks
if [ "${1:-}" = build ]; then
  rm -rf $V; mkdir -p $V
  for v in $VARIANTS; do
    IFS=: read b m k <<< "$v"
    nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -lineinfo -std=c++17 -Xcompiler -fPIC -shared \
      -DPT_NODE_BURST=$b -DPT_MIN_NODE_LANES=$m -DPT_TRAV_MIN_BLOCKS=$k $HERE/path-tracing_b200/csrc/pt_api.cu -o $V/libptrender_b${b}_m${m}_k${k}.so &
  done
  wait; ls -la $V
else
  mkdir -p $HERE/gpurun_out
  for v in $VARIANTS; do
    IFS=: read b m k <<< "$v"
    PTB200_RENDER_SO=$V/libptrender_b${b}_m${m}_k${k}.so python $HERE/bench.py --workload torus1m --spp 128 --steps 2 --warmup 1 --no-cpu \
      2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('b$b m$m k$k', round(d['value']), 'Mrays/s', {k: round(v,2) for k,v in r['class_ms'].items()}, 'e2e calls', [c['ms'] for c in d['e2e']['calls']])"
  done | tee $HERE/gpurun_out/sweep_traverse.txt
fi
