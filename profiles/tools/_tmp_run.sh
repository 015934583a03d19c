CMD="python bench.py --steps 2 --warmup 3 --no-graph --no-cpu-baseline --no-ref-cuda"
$CMD > gpurun_out/r2r_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 100 -c 300 --csv --log-file gpurun_out/r2r_launches.csv $CMD > gpurun_out/r2r_ncu1.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"blend_backward|blend_forward|gauss_backward|preprocess_kernel|tile_scatter|tile_count|sort_pass" -s 14 -c 11 -o gpurun_out/r2r_full $CMD > gpurun_out/r2r_ncu2.log 2>&1
echo "full rc=$?"; ls -la gpurun_out/r2r_*
