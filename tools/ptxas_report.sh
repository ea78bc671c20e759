#!/usr/bin/env bash
# Registers / spills / smem of every kernel in one .cu:  tools/ptxas_report.sh dgs_forward.cu [filter]
cd "$(dirname "$0")/../diff_gaussian_sampling_b200/csrc"
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr --extended-lambda -Xptxas -v -c "$1" -o /tmp/ptxas_report.o 2>&1 \
 | awk '/Compiling entry function/ {name=$7} /bytes stack frame/ {spill=$0} /Used [0-9]+ registers/ {print name, $0, "|", spill}' \
 | sed -E "s/ptxas info    : //g; s/Function properties for [^ ]+//" | c++filt | grep -E "${2:-.}"
