#!/bin/bash
# bench.py under torchrun on N GPUs of one box (the driver's launch line): tools/gpu_scale.sh N tag
n=$1; tag=${2:-r02c}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n \
  > gpurun_out/${tag}_bench_small256_n$n.json 2> gpurun_out/${tag}_bench_small256_n$n.err
echo "rc=$?"
tail -c 3000 gpurun_out/${tag}_bench_small256_n$n.json | cut -c1-1500
tail -3 gpurun_out/${tag}_bench_small256_n$n.err
