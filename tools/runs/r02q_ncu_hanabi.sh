# ncu evidence for the round-2 Hanabi kernels (register-resident records, one CTA per SM), 65 536 worlds, 1 GPU.
# Each capture only after the same command has exited 0 without ncu.  STEPS PER LAUNCH: fused = 200, per-step = 1.
mkdir -p gpurun_out
python tools/debug/hanabi_roll_once.py 200 || exit 1
python tools/debug/hanabi_roll_once.py 200 per_step || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r02q_hanabi_launches.csv python tools/debug/hanabi_roll_once.py 20 per_step > gpurun_out/r02q_ncu0.log 2>&1; echo launches rc=$?
ncu --set full --clock-control none --import-source on -k regex:rollout_kernel -s 2 -c 1 -f -o gpurun_out/r02q_hanabi_rollout python tools/debug/hanabi_roll_once.py 200 > gpurun_out/r02q_ncu_roll.log 2>&1; echo rc=$?
ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 300 -c 1 -f -o gpurun_out/r02q_hanabi_step python tools/debug/hanabi_roll_once.py 200 per_step > gpurun_out/r02q_ncu_step.log 2>&1; echo rc=$?
ls -la gpurun_out/*.ncu-rep
