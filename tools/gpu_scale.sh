#!/bin/bash
# Scaling session on an N-GPU box (gpurun --gpus 8 -- 'bash tools/gpu_scale.sh'): bench.py at 1, 2, 4, 8 ranks the way
# the driver launches it (torchrun, 127.0.0.1), Cornell at every N, the 1M-triangle scene and the 4K composite at the
# ends.  Lines go to gpurun_out/scale_*.json.
set -u
mkdir -p gpurun_out
O=gpurun_out
MAXN=${1:-8}
nvidia-smi --query-gpu=index,name,clocks.sm --format=csv > $O/scale_smi.txt 2>&1
run() {  # workload n
  local w=$1 n=$2
  if [ "$n" = 1 ]; then
    python bench.py --workload $w --gpus 1 --no-cpu > $O/scale_${w}_n$n.json 2> $O/scale_${w}_n$n.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --workload $w --gpus $n --no-cpu > $O/scale_${w}_n$n.json 2> $O/scale_${w}_n$n.err
  fi
  echo "rc=$?" >> $O/scale_${w}_n$n.err
}
for n in 2 4 8; do [ $n -le $MAXN ] && run cornell $n; done
run torus1m $MAXN
run composite4k $MAXN
ls -la $O
