#!/bin/bash
# one B200: launch list + full capture (with source) of the current render kernel on C3, one frame at a time; host topology
O=gpurun_out
(lscpu | grep -i 'numa\|socket\|model name\|^CPU(s)'; nvidia-smi topo -m; for d in /sys/bus/pci/devices/*; do if [ -f $d/numa_node ] && grep -qi 0x10de $d/vendor 2>/dev/null; then echo "$d $(cat $d/numa_node) $(cat $d/class)"; fi; done; ls /sys/devices/system/node/ 2>/dev/null; cat /sys/devices/system/node/node*/cpulist 2>/dev/null; nproc; cat /proc/self/status | grep -i 'cpus_allowed_list\|mems_allowed_list') > $O/r02m_topology.txt 2>&1
B="python bench.py --steps 2 --warmup 3 --cpu-seconds 0 --in-flight 1"
$B > $O/r02m_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02m_launches_march_c3.csv $B > $O/r02m_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:renderKernel -s 3 -c 1 -f -o $O/r02m_prof_march $B > $O/r02m_ncu_full.log 2>&1
ls -la $O/*.ncu-rep; tail -3 $O/r02m_ncu_full.log; cat $O/r02m_topology.txt | head -60
