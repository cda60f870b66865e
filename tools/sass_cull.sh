#!/bin/bash
# Offline look at cull_kernel's SASS: static instruction count, registers, opcode histogram; listing in /tmp/cull.lst
set -e
cd "$(dirname "$0")/.."
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -prec-div=true -prec-sqrt=true "$@" \
  -Xptxas -v -cubin -o /tmp/cull.cubin ravengine_b200/csrc/cull.cu 2>&1 | grep -A1 "cull_kernel" | grep -E "registers|spill" || true
cuobjdump -sass /tmp/cull.cubin | awk '/Function : /{f=($0 ~ /cull_kernel/)} f' | grep -E "^\s+/\*[0-9a-f]{4}\*/" | sed -E 's#^\s+/\*([0-9a-f]{4})\*/\s+(.*);.*#\1 \2#' > /tmp/cull.lst
echo "static instructions: $(grep -vc ' NOP' /tmp/cull.lst)"
awk '{op=$2; if (op ~ /^@/) op=$3; sub(/\..*/,"",op); c[op]++} END{for(k in c) print c[k], k}' /tmp/cull.lst | sort -rn | head -${TOP:-14} | tr '\n' ';'; echo
