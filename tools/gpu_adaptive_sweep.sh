#!/bin/bash
# Samples per wavefront step under the ADAPTIVE rule (error > 0): frame time and traced / committed segments.
set -u
O=gpurun_out; mkdir -p $O
line() { python -c "import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('$1', round(d['value']), 'Mrays/s', round(d['ms_per_step'],2), 'ms launches', d['gpu_launches'], 'segments', int(d['config']['segments_per_frame']), {k: round(v,2) for k,v in r['class_ms'].items()})"; }
for q in 2 3 4 5 6; do
  python bench.py --workload tor --no-cpu --q-shift $q --spp 256 --steps 2 --warmup 1 2>/dev/null | line "tor512 256spp err1e-3 q$q"
  python bench.py --workload cornell --no-cpu --q-shift $q --error 0.001 --steps 2 --warmup 1 2>/dev/null | line "cornell1080p 256spp err1e-3 q$q"
  python bench.py --workload cornell --no-cpu --q-shift $q --error 0.02 --steps 2 --warmup 1 2>/dev/null | line "cornell1080p 256spp err2e-2 q$q"
done | tee $O/sweep_adaptive.txt
