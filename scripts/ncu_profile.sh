# ncu evidence for one round: launch list of every product kernel + one full capture per hot kernel.
# usage (on the GPU box): bash scripts/ncu_profile.sh <tag>
set -x
TAG=${1:-rX}
CMD="python bench.py --no-f32 --mono-seqs 60 --grf-seqs 64 --shutter-seqs 0 --seqs 64 --wave 32 --lm-seqs 16 --lm-iters 4 --cpu-seqs 0 --steps 2 --warmup 1 --no-e2e"
K="eval_kernel|assemble_kernel|band_|step_apply|reduce_|step_accept|update_multipliers|prior_kernel|eom_|grf_"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"$K" -c 1500 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"eval_kernel" -s 1 -c 2 -o gpurun_out/${TAG}_evaljac $CMD > gpurun_out/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"assemble_kernel" -c 1 -o gpurun_out/${TAG}_assemble $CMD > gpurun_out/ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"band_solve" -c 1 -o gpurun_out/${TAG}_bandsolve $CMD > gpurun_out/ncu4.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"band_block" -s 3 -c 1 -o gpurun_out/${TAG}_bandblock $CMD > gpurun_out/ncu5.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"grf_solve" -s 1 -c 1 -o gpurun_out/${TAG}_grf $CMD > gpurun_out/ncu6.log 2>&1
cp cheetah_pose_estimation_b200/libcpe_b200.so gpurun_out/${TAG}_libcpe_b200.so; ls -la gpurun_out
