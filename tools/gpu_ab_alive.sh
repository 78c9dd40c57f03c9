#!/bin/bash
# A/B of render-library variants over dense, Russian-roulette and adaptive workloads (run on the GPU box)
set -u
HERE=$(cd "$(dirname "$0")/.." && pwd); V=$HERE/path-tracing_b200/lib/variants; O=$HERE/gpurun_out; mkdir -p $O
line() { python -c "import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('$1', round(d['value']), 'Mrays/s', round(d['ms_per_step'],2), 'ms', {k: round(v,2) for k,v in r['class_ms'].items()})"; }
for name in ${VARIANTS:-dense alive}; do
  export PTB200_RENDER_SO=$V/libptrender_$name.so
  echo "$name parity: $(python -m pytest $HERE/tests -m gpu -x -q 2>&1 | tail -1)"
  python bench.py --workload cornell --no-cpu 2>/dev/null | line "$name cornell flat"
  python bench.py --workload torus1m --no-cpu 2>/dev/null | line "$name torus1m flat"
  python bench.py --workload glossy_rr --no-cpu --steps 2 --warmup 1 2>/dev/null | line "$name glossy_rr"
  python bench.py --workload tor --no-cpu 2>/dev/null | line "$name tor C1 adaptive"
  python bench.py --workload cornell --no-cpu --error 0.001 --steps 2 --warmup 1 2>/dev/null | line "$name cornell adaptive 1e-3"
done | tee $O/ab_alive.txt
