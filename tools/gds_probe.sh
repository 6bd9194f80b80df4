#!/bin/bash
# GPUDirect Storage: is there anything to use on this box?  ("probe, don't assume")
echo "nvidia_fs module: $(lsmod 2>/dev/null | grep -c nvidia_fs) loaded; /proc/driver/nvidia-fs: $(ls /proc/driver/nvidia-fs 2>/dev/null | wc -l) entries"
echo "libcufile: $(ls /usr/local/cuda/lib64/libcufile.so* 2>/dev/null | head -1)"
echo "gdscheck: $(ls /usr/local/cuda/gds/tools/gdscheck* 2>/dev/null | head -1)"
if [ -x /usr/local/cuda/gds/tools/gdscheck ]; then timeout 30 /usr/local/cuda/gds/tools/gdscheck -p 2>&1 | head -40; fi
echo "file systems:"; df -T /dev/shm /tmp / 2>/dev/null
echo "block devices:"; lsblk -d -o NAME,SIZE,ROTA,MODEL 2>/dev/null | head
