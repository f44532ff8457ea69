#!/bin/bash
# builds experimental variants of the library (fp32 'I' kernel only) into tools/variants/<name>.so
# usage: tools/variants.sh name "-DFLR_MAXT=224 -DFLR_MINB=3 ..."
set -e
cd "$(dirname "$0")/.."
name=$1; flags=$2
mkdir -p build/var_$name tools/variants
cd pregenerating-pde_b200/csrc
for f in capi bias_act upfirdn2d flrelu_act cin flrelu_fused flrelu_fused_f32 flrelu_plane_f32; do
  nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -DFLR_QUICK $flags -c $f.cu -o ../../build/var_$name/$f.o &
done
wait
cd ../..
nvcc -shared -o tools/variants/$name.so build/var_$name/*.o -gencode arch=compute_100a,code=sm_100a --cudart static
cuobjdump --dump-resource-usage build/var_$name/flrelu_fused_f32.o 2>/dev/null | grep -A1 "Function" | paste - - | sed -E 's/.*Cfg(ILi[0-9]ELi[0-9]ELi[0-9]ELi[0-9]EEELi[0-9]).*REG:([0-9]+) STACK:([0-9]+).*/\1 REG \2 STACK \3/' | sort | sed "s/^/$name: /"
