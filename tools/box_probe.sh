#!/bin/bash
# What the GPU box looks like from the host side (memory, cores, NUMA, PCIe topology) + the file->device probe.
mkdir -p gpurun_out
O=gpurun_out/box_probe.txt
{
  echo "== nproc"; nproc
  echo "== free -g"; free -g
  echo "== df"; df -h /dev/shm /tmp . 2>&1
  echo "== lscpu"; lscpu | egrep -i 'model name|socket|numa|thread|core|L3' 
  echo "== numa nodes"; ls /sys/devices/system/node/ 2>&1 | head; cat /sys/devices/system/node/node*/meminfo 2>/dev/null | egrep 'MemTotal|MemFree'
  echo "== nvidia-smi -L"; nvidia-smi -L
  echo "== topo"; nvidia-smi topo -m 2>&1
  echo "== pcie"; nvidia-smi --query-gpu=index,pci.bus_id,pcie.link.gen.current,pcie.link.width.current,pcie.link.gen.max --format=csv 2>&1
  echo "== ulimit -l"; ulimit -l
  echo "== THP"; cat /sys/kernel/mm/transparent_hugepage/enabled /sys/kernel/mm/transparent_hugepage/shmem_enabled 2>&1
} > $O 2>&1
./tools/hostreg_probe ${1:-4} >> $O 2>&1
cat $O
