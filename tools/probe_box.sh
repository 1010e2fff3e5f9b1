#!/bin/bash
# facts about the GPU box that size the transfer pipeline (run under gpurun)
echo "== cpu"; nproc; lscpu | grep -E "Model name|Socket|NUMA|Thread|Core" ; free -g | head -2
echo "== numa"; ls /sys/devices/system/node/ | grep node; cat /sys/devices/system/node/node*/cpulist 2>/dev/null
echo "== disks"; df -h / /tmp /dev/shm 2>/dev/null; mount | grep -E " / | /tmp | /dev/shm " 
echo "== gpu"; nvidia-smi --query-gpu=index,name,pci.bus_id,pcie.link.gen.current,pcie.link.width.current,memory.total --format=csv
nvidia-smi topo -m 2>/dev/null | head -20
echo "== gds"; ls /usr/local/cuda/gds/tools 2>/dev/null | head; lsmod 2>/dev/null | grep -E "nvidia_fs|nvidia" | head
echo "== kernel"; uname -r
