mkdir -p gpurun_out
python tools/debug/hanabi_roll_once.py 200 || exit 1
ncu --set full --clock-control none --import-source on -k regex:rollout_kernel -s 2 -c 1 -f -o gpurun_out/r02k_hanabi_rollout python tools/debug/hanabi_roll_once.py 200 > gpurun_out/r02k_ncu_roll.log 2>&1; echo rc=$?
ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 300 -c 1 -f -o gpurun_out/r02k_hanabi_step python tools/debug/hanabi_roll_once.py 200 per_step > gpurun_out/r02k_ncu_step.log 2>&1; echo rc=$?
ls -la gpurun_out/*.ncu-rep
