#!/bin/bash
# NUMA placement of the GPUs and its effect on pinned-memory copy rates (bench.py e2e ceiling)
lscpu | grep -i "numa\|^CPU(s)\|Model name" 
nvidia-smi topo -m 2>/dev/null | head -20
for d in /sys/bus/pci/devices/*; do
  if [ -f $d/class ] && grep -q "^0x0302\|^0x0300" $d/class 2>/dev/null; then echo "$d numa_node=$(cat $d/numa_node) local_cpulist=$(cat $d/local_cpulist)"; fi
done
python - <<'PY'
import os
print("affinity now:", len(os.sched_getaffinity(0)), sorted(os.sched_getaffinity(0))[:8], "...")
PY
which numactl taskset
python scripts/pcie_probe.py
