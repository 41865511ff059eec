set -x
CMD="python bench.py --seqs 64 --wave 32 --lm-seqs 16 --lm-iters 4 --cpu-seqs 0 --steps 2 --warmup 1 --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"frame_kernel|band_solve|step_apply|reduce_cost|step_accept|update_multipliers" -c 400 --csv --log-file gpurun_out/r1a_launches.csv $CMD > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"frame_kernel<1>" -s 1 -c 2 -o gpurun_out/r1a_evaljac $CMD > gpurun_out/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"frame_kernel<2>" -c 1 -o gpurun_out/r1a_assemble $CMD > gpurun_out/ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"band_solve" -c 1 -o gpurun_out/r1a_bandsolve $CMD > gpurun_out/ncu4.log 2>&1
tail -3 gpurun_out/ncu*.log
ls -la gpurun_out
