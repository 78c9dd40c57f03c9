#!/bin/bash
# A/B harness for compile-time variants of the render library.
#   VARIANTS="name:-DFOO=1,-DBAR=2 other:" tools/ab_variants.sh build     (here: nvcc cross-compiles)
#   VARIANTS=... WORKLOADS="torus1m tor" tools/ab_variants.sh run         (on the GPU box, via gpurun)
# Each variant is checked against the oracle-pinned goldens first (a variant that fails parity prints FAIL).
set -u
HERE=$(cd "$(dirname "$0")/.." && pwd)
V=$HERE/path-tracing_b200/lib/variants
VARIANTS="${VARIANTS:-base:}"
WORKLOADS="${WORKLOADS:-torus1m}"
if [ "${1:-}" = build ]; then
  rm -rf $V; mkdir -p $V
  for v in $VARIANTS; do
    name=${v%%:*}; flags=${v#*:}; flags=${flags//,/ }
    nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -lineinfo -std=c++17 -Xcompiler -fPIC -shared \
      $flags $HERE/path-tracing_b200/csrc/pt_api.cu -o $V/libptrender_$name.so -Xptxas -v 2> $V/$name.ptxas &
  done
  wait; ls -la $V
  for v in $VARIANTS; do name=${v%%:*}; echo "$name: $(grep -A1 'k_wf_traverseILb0' $V/$name.ptxas | grep -o 'Used [0-9]* registers' | head -1) $(grep -A1 'k_wf_traverseILb0' $V/$name.ptxas | grep -o '[0-9]* bytes stack frame, [0-9]* bytes spill stores' | head -1)"; done
else
  mkdir -p $HERE/gpurun_out
  for v in $VARIANTS; do
    name=${v%%:*}
    export PTB200_RENDER_SO=$V/libptrender_$name.so
    ok=$(python -m pytest $HERE/tests -m gpu -x -q -k "${PARITY:-nearest_hit_golden or render_golden or lbvh_equals_brute}" 2>&1 | tail -1)
    for w in $WORKLOADS; do
      python $HERE/bench.py --workload $w --steps 3 --warmup 2 --no-cpu 2>/dev/null | \
        python -c "import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('$name $w', round(d['value']), 'Mrays/s', {k: round(v,2) for k,v in r['class_ms'].items()}, '| parity: $ok')"
    done
  done | tee $HERE/gpurun_out/ab_variants.txt
fi
