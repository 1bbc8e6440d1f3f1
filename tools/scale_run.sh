#!/bin/bash
# Runs bench.py at N = 1, 2, 4, 8 on one box (as the driver does) and writes gpurun_out/scale_N.json.
# e2e at N > 2 pins N x 10 GB of host memory: only attempted when the box has room for it.
mkdir -p gpurun_out
nvidia-smi -L | wc -l > gpurun_out/scale_gpus.txt
free -g | tee -a gpurun_out/scale_gpus.txt
avail=$(awk '/MemAvailable/ {print int($2/1048576)}' /proc/meminfo)
maxn=${1:-8}
for n in 1 2 4 8; do
  [ $n -gt $maxn ] && break
  extra=""
  if [ $((n * 12 + 20)) -gt $avail ]; then extra="--no-e2e"; fi
  if [ $n -eq 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 30 --warmup 5 $extra > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --gpus $n --steps 30 --warmup 5 $extra > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
  fi
  echo "N=$n rc=$? $(cut -c1-180 gpurun_out/scale_$n.json)"
done
