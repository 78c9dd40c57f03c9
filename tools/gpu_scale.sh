#!/bin/bash
# Scaling sessions.  Box time is charged N-fold, so every width runs on a box of its own size:
#   gpurun --gpus 2 -- 'bash tools/gpu_scale.sh 2 tests torus1m reference'
#   gpurun --gpus 4 -- 'bash tools/gpu_scale.sh 4 torus1m'
#   gpurun --gpus 8 -- 'bash tools/gpu_scale.sh 8 tests torus1m composite4k'
# bench.py is launched the way the driver launches it (torchrun, 127.0.0.1); `tests` = the multi-GPU parity tests
# (in-process NCCL path at every width the box has); `reference` = the reference arm under torchrun (its `cores`
# must be the box's).  Lines go to gpurun_out/scale_*.json.
set -u
mkdir -p gpurun_out
O=gpurun_out
N=${1:-8}; shift
nvidia-smi --query-gpu=index,name,clocks.sm --format=csv > $O/scale_smi_n$N.txt 2>&1
for what in "$@"; do
  case $what in
    tests)
      ( time python -m pytest tests/test_gpu_parity.py -m gpu -q -k "in_process_multi_gpu or row_bands or scene_cache" ) \
        > $O/scale_pytest_multi_gpu_n$N.log 2>&1; echo "pytest rc=$?" >> $O/scale_pytest_multi_gpu_n$N.log ;;
    reference)
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29600 \
        bench.py --impl reference --gpus $N --steps 1 --warmup 0 > $O/scale_reference_torus1m_n$N.json 2> $O/scale_reference_n$N.err ;;
    *)
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) \
        bench.py --workload $what --gpus $N --no-cpu > $O/scale_${what}_n$N.json 2> $O/scale_${what}_n$N.err
      echo "rc=$?" >> $O/scale_${what}_n$N.err ;;
  esac
done
ls -la $O
