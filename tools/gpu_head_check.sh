#!/bin/bash
# Short check of the tree as it stands on one B200: GPU parity tests, smoke, the driver's two bench arms.
set -u
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
( time python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
( time python bench.py ) > $O/bench_torus1m.json 2> $O/bench_torus1m.err; echo "bench torus1m rc=$?" >> $O/bench_torus1m.err
( time python bench.py --impl reference ) > $O/bench_reference_torus1m.json 2> $O/bench_reference.err
python bench.py --workload cornell --no-cpu > $O/bench_cornell.json 2> $O/bench_cornell.err; echo "bench cornell rc=$?" >> $O/bench_cornell.err
tail -3 $O/pytest_gpu.log; cat $O/smoke.log; tail -c 600 $O/bench_torus1m.json; tail -4 $O/bench_torus1m.err
