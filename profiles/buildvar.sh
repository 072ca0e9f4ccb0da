#!/bin/bash
# usage: buildvar.sh name -Dflags...   -> slam_b200/lib/<name>.so
name=$1; shift
cd /root/repo
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC -shared -cudart static "$@" -o slam_b200/lib/$name.so slam_b200/csrc/ccicp_api.cu slam_b200/csrc/ccicp_multi.cu -ldl 2>&1 | grep -v "^$" | head -20
ls -la slam_b200/lib/$name.so
