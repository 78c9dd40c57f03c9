#!/bin/bash
# Round-2 evidence session on one B200 (run through gpurun): parity tests, smoke, bench lines of every workload, the
# reference arm, the watertight report, then the ncu passes (launch list + full captures of production-shaped launches).
# Numbers printed under ncu are never bench values.  gpurun_out/ must stay below 64 MiB: the .ncu-rep files are
# exported to CSV on the box and removed.
set -u
mkdir -p gpurun_out
O=gpurun_out
WHAT=${1:-all}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
( time python -m pytest tests -m gpu -q ) > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
python bench.py > $O/bench_torus1m.json 2> $O/bench_torus1m.err; echo "bench torus1m rc=$?" >> $O/bench_torus1m.err
for w in cornell tor composite4k glossy glossy_rr; do
  python bench.py --workload $w > $O/bench_$w.json 2> $O/bench_$w.err; echo "bench $w rc=$?" >> $O/bench_$w.err
done
python bench.py --impl reference > $O/bench_reference_torus1m.json 2> $O/bench_reference.err
python bench.py --impl reference --workload cornell > $O/bench_reference_cornell.json 2>> $O/bench_reference.err
python tools/watertight_report.py > $O/watertight.log 2>&1
python tools/bvh_stats.py > $O/bvh_stats.log 2>&1
[ "$WHAT" = "noprof" ] && exit 0
python bench.py --spp 64 --steps 1 --warmup 1 --no-cpu > $O/plain_launches.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_torus1m.csv \
    python bench.py --spp 64 --steps 1 --warmup 1 --no-cpu > $O/ncu_launches.log 2>&1
for w in torus1m cornell; do
  python tools/prof_one.py $w 64 > $O/plain_$w.log 2>&1 && \
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_ -s 1 -c 18 -f -o /tmp/full_$w \
      python tools/prof_one.py $w 64 > $O/ncu_full_$w.log 2>&1
  ncu -i /tmp/full_$w.ncu-rep --page raw --csv > $O/full_${w}_raw.csv 2>/dev/null
  ncu -i /tmp/full_$w.ncu-rep --page source --csv --print-source sass > $O/full_${w}_sass.csv 2>/dev/null
  gzip -f $O/full_${w}_sass.csv
done
du -sh $O; ls -la $O
