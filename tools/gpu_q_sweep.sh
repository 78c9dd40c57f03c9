#!/bin/bash
# Samples-per-step sweep on both bench workloads + one ncu capture of the traversal kernel (run via gpurun).
set -u
O=gpurun_out; mkdir -p $O
( python -m pytest tests -m gpu -x -q 2>&1 | tail -3 ) > $O/pytest_gpu.log
for w in torus1m cornell; do
  for q in 2 4 5 6; do
    python bench.py --workload $w --spp 64 --steps 2 --warmup 1 --no-cpu --q-shift $q 2>/dev/null | \
      python -c "import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('$w q$q', round(d['value']), 'Mrays/s e2e', round(d['e2e']['value']), {k: round(v,2) for k,v in r['class_ms'].items()})"
  done
done | tee $O/sweep_q.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_wf_traverse --launch-skip 10 -c 2 -f -o /tmp/full_trav \
  python bench.py --workload torus1m --spp 16 --steps 1 --warmup 1 --no-cpu --q-shift 4 > $O/ncu_full_trav.log 2>&1
ncu -i /tmp/full_trav.ncu-rep --page raw --csv > $O/full_trav_raw.csv 2>/dev/null
ncu -i /tmp/full_trav.ncu-rep --page source --csv --print-source sass > $O/full_trav_sass.csv 2>/dev/null
gzip -f $O/full_trav_sass.csv
cat $O/pytest_gpu.log
