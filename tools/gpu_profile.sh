#!/bin/bash
# ncu passes of the tree as it stands (one B200): launch list of the bench command, then `--set full` captures of one
# production-shaped step (1920x1080 x 64 samples) per workload, exported to CSV on the box (the .ncu-rep files are too
# large for gpurun_out).  Numbers printed under ncu are never bench values.
set -u
mkdir -p gpurun_out
O=gpurun_out
python bench.py --spp 64 --steps 1 --warmup 1 --no-cpu > $O/plain_launches.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_torus1m.csv \
    python bench.py --spp 64 --steps 1 --warmup 1 --no-cpu > $O/ncu_launches.log 2>&1
for w in "$@"; do
  python tools/prof_one.py $w 64 > $O/plain_$w.log 2>&1 && \
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_ -s 1 -c 18 -f -o /tmp/full_$w \
      python tools/prof_one.py $w 64 > $O/ncu_full_$w.log 2>&1
  ncu -i /tmp/full_$w.ncu-rep --page raw --csv > $O/full_${w}_raw.csv 2>/dev/null
  ncu -i /tmp/full_$w.ncu-rep --page source --csv --print-source sass > $O/full_${w}_sass.csv 2>/dev/null
  ncu -i /tmp/full_$w.ncu-rep --page source --csv --print-source cuda > $O/full_${w}_cuda.csv 2>/dev/null
  gzip -f $O/full_${w}_sass.csv $O/full_${w}_cuda.csv
done
du -sh $O; ls -la $O
