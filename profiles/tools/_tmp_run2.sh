#!/bin/bash
cd /root/repo
for v in p5; do
RG_LIB=/root/repo/riggs_b200/lib/libriggs_b200_$v.so timeout 120 python profiles/tools/mlp_prof.py > gpurun_out/mlp_prof_$v.txt 2>&1
echo $v; sed -n 3,3p gpurun_out/mlp_prof_$v.txt; sed -n 5,5p gpurun_out/mlp_prof_$v.txt; tail -n 1 gpurun_out/mlp_prof_$v.txt
done
