#!/bin/bash
# multi-GPU evidence on an N-GPU box: the 2-GPU tests, the default bench, BASELINE configs[3] (one 100 GB file streamed,
# split across N ranks, and the same through ONE process driving N devices), and at N = 8 the configs[4] sweep
N=${1:-2}
mkdir -p gpurun_out
O=gpurun_out
{ nvidia-smi -L; nvidia-smi topo -m; free -g; nproc; lscpu | egrep -i 'socket|numa'; } > $O/r2_box_n$N.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ "$2" != "notests" ]; then
  timeout 600 python -m pytest tests/test_multi_gpu.py tests/test_cpp_api.py -m gpu -x -q -k "multi or exchange or peer or local" > $O/r2_pytest_n$N.log 2>&1
  echo "pytest rc=$?"; tail -3 $O/r2_pytest_n$N.log
fi
timeout 900 $TR --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > $O/r2_scale_$N.json 2> $O/r2_scale_$N.err; echo "scale rc=$?"; tail -2 $O/r2_scale_$N.err; cut -c1-200 $O/r2_scale_$N.json
W=line_byte_offsets+lines/needle32/100GB-streamed
timeout 1200 $TR --master-port 29512 bench.py --gpus $N --workload $W --steps 10 --warmup 3 --e2e-steps 3 > $O/r2_cfg3_n$N.json 2> $O/r2_cfg3_n$N.err; echo "cfg3 rc=$?"; tail -2 $O/r2_cfg3_n$N.err; cut -c1-200 $O/r2_cfg3_n$N.json
timeout 1200 python bench.py --gpus $N --single-process --workload $W --warmup 2 --e2e-steps 3 > $O/r2_cfg3_sp_n$N.json 2> $O/r2_cfg3_sp_n$N.err; echo "cfg3 single-process rc=$?"; tail -2 $O/r2_cfg3_sp_n$N.err; cut -c1-200 $O/r2_cfg3_sp_n$N.json
if [ $N -eq 8 ]; then
  : > $O/r2_sweep8_400gb.jsonl
  for m in 1 2 3 4 8 16 32 64; do
    timeout 600 $TR --master-port $((29520 + m)) bench.py --gpus 8 --workload count/sweep/50GB --pattern-len $m --steps 10 --warmup 3 --no-e2e \
      2> $O/r2_sweep8_$m.err | tail -1 >> $O/r2_sweep8_400gb.jsonl
    echo "sweep m=$m rc=$?"; tail -1 $O/r2_sweep8_400gb.jsonl | cut -c1-160
  done
fi
