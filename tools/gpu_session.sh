#!/bin/bash
# One GPU-box session: parity tests, smoke, bench lines, ncu launch list + one full capture per workload.
# Run via gpurun.  Numbers printed under ncu are never bench values.  gpurun_out/ must stay below 64 MiB, so the
# .ncu-rep files are exported to CSV on the box and removed.
set -u
mkdir -p gpurun_out
O=gpurun_out
WHAT=${1:-all}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
( time python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
for w in cornell torus1m tor glossy; do
  python bench.py --workload $w > $O/bench_$w.json 2> $O/bench_$w.err; echo "bench $w rc=$?" >> $O/bench_$w.err
done
python bench.py --impl reference > $O/bench_reference.json 2> $O/bench_reference.err
[ "$WHAT" = "noprof" ] && exit 0
for w in cornell torus1m; do
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_$w.csv \
    python bench.py --workload $w --spp 8 --steps 1 --warmup 1 --no-cpu > $O/ncu_launches_$w.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_ --launch-skip 20 -c 6 -f -o /tmp/full_$w \
    python bench.py --workload $w --spp 4 --steps 1 --warmup 1 --no-cpu > $O/ncu_full_$w.log 2>&1
  ncu -i /tmp/full_$w.ncu-rep --page raw --csv > $O/full_${w}_raw.csv 2>/dev/null
  ncu -i /tmp/full_$w.ncu-rep --page source --csv --print-source sass > $O/full_${w}_sass.csv 2>/dev/null
  ncu -i /tmp/full_$w.ncu-rep --page details > $O/full_${w}_details.txt 2>/dev/null
  gzip -f $O/full_${w}_sass.csv
done
du -sh $O; ls -la $O
