# ncu evidence of round 2 for the bench command (1 GPU).  Kernels run one at a time under ncu: BENCH_NO_GATE=1.
# Each capture only after the same command has exited 0 without ncu.
mkdir -p gpurun_out
CMD="python bench.py --steps 400 --warmup 5 --no-extra --no-cpu-baseline"
$CMD > gpurun_out/r02e_bench400.json 2> gpurun_out/r02e_bench400.err || { echo "bench failed"; tail -5 gpurun_out/r02e_bench400.err; exit 1; }
export BENCH_NO_GATE=1
# launch list: the first 300 launches of the command (warm-up, the sizing region, then timed 400-step fused launches)
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r02e_launches.csv $CMD > gpurun_out/r02e_ncu1.log 2>&1; echo launches rc=$?
# one --set full capture of the dominant kernel: the third 400-step rollout_tile launch
ncu --set full --clock-control none --import-source on -k regex:rollout_tile -s 2 -c 1 -f -o gpurun_out/r02e_rollout_tile $CMD > gpurun_out/r02e_ncu2.log 2>&1; echo full rc=$?
# ... and of the per-step kernel (what env.step() runs): one launch in steady state
ncu --set full --clock-control none --import-source on -k regex:step_tile -s 50 -c 1 -f -o gpurun_out/r02e_step_tile $CMD > gpurun_out/r02e_ncu3.log 2>&1; echo full2 rc=$?
ls -la gpurun_out/
