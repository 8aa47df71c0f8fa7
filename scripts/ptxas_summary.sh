#!/bin/bash
# registers / spills of the hot kernels as ptxas reports them (no GPU needed): scripts/ptxas_summary.sh [extra nvcc flags]
cd "$(dirname "$0")/../opencl_pathtracer_b200"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -fmad=false -std=c++17 -Xcompiler -fPIC,-fopenmp,-ffp-contract=off \
  -I ../include -I csrc -Xptxas -v "$@" -c csrc/pt_kernels.cu -o /tmp/ptk_summary.o 2>&1 | python3 -c '
import re, sys, subprocess
txt = sys.stdin.read()
for m in re.finditer(r"Compiling entry function .(\S+?). for.*?\n.*?\n\s*(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\n.*?Used (\d+) registers", txt):
    name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
    if "k_trace" in name or "k_shade" in name or "k_sort" in name:
        print("%-70s regs %3s stack %3s spill st/ld %3s/%3s" % (name.replace("void ptc::", ""), m.group(5), m.group(2), m.group(3), m.group(4)))
if "error" in txt: print(txt)
'
