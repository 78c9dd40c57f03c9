#!/bin/bash
# Scaling session on an N-GPU box (gpurun --gpus 8 -- 'bash tools/gpu_scale.sh 8'): the multi-GPU parity tests, then
# bench.py on BASELINE configs[2] (torus1m) at 1, 2, 4, 8 ranks the way the driver launches it (torchrun, 127.0.0.1),
# the 4K composite and Cornell at the full width, and the reference arm under torchrun (its `cores` must be the box's).
# Lines go to gpurun_out/scale_*.json.  Box time is charged N-fold: nothing here runs longer than it has to.
set -u
mkdir -p gpurun_out
O=gpurun_out
MAXN=${1:-8}
nvidia-smi --query-gpu=index,name,clocks.sm --format=csv > $O/scale_smi.txt 2>&1
( time python -m pytest tests/test_gpu_parity.py -m gpu -q -k "in_process_multi_gpu or row_bands or scene_cache" ) \
  > $O/scale_pytest_multi_gpu.log 2>&1; echo "pytest rc=$?" >> $O/scale_pytest_multi_gpu.log
run() {  # workload n [extra bench.py flags]
  local w=$1 n=$2; shift 2
  if [ "$n" = 1 ]; then
    python bench.py --workload $w --gpus 1 --no-cpu "$@" > $O/scale_${w}_n$n.json 2> $O/scale_${w}_n$n.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --workload $w --gpus $n --no-cpu "$@" > $O/scale_${w}_n$n.json 2> $O/scale_${w}_n$n.err
  fi
  echo "rc=$?" >> $O/scale_${w}_n$n.err
}
for n in 1 2 4 8; do [ $n -le $MAXN ] && run torus1m $n; done
run composite4k $MAXN
run cornell $MAXN
python -m torch.distributed.run --nnodes=1 --nproc-per-node $MAXN --master-addr 127.0.0.1 --master-port 29600 \
  bench.py --impl reference --gpus $MAXN --steps 1 --warmup 0 > $O/scale_reference_torus1m_n$MAXN.json 2> $O/scale_reference.err
ls -la $O
