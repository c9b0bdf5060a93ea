#!/bin/bash
# Dev tool: bench.py on N GPUs of one box the way the driver launches it; usage: gpu_multi.sh N tag [bench args]
N=$1; TAG=$2; shift 2
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus $N "$@" > gpurun_out/${TAG}.json 2> gpurun_out/${TAG}.err
echo "rc=$? $(wc -c < gpurun_out/${TAG}.json) bytes"; tail -c 1500 gpurun_out/${TAG}.err
