# 8-GPU lease of round 2: the driver's own bench command at 8 ranks, BASELINE config 5 at 8 x 8 192 worlds, and the
# device->host ceiling of the box.  Run as: gpurun --gpus 8 -- 'bash tools/runs/r02c_8gpu.sh'  (outputs under gpurun_out/)
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r02c_topo.txt 2>&1; nproc >> gpurun_out/r02c_topo.txt; lscpu | head -25 >> gpurun_out/r02c_topo.txt
$TR --nproc-per-node 8 --master-port 29501 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r02c_bench_8gpu.json 2> gpurun_out/r02c_bench_8gpu.err; echo bench8 rc=$?
$TR --nproc-per-node 8 --master-port 29502 bench.py --gpus 8 --steps 20 --warmup 5 --workload many_player --no-cpu-baseline > gpurun_out/r02c_many_player_8gpu.json 2> gpurun_out/r02c_many_player_8gpu.err; echo many_player8 rc=$?
for n in 8 4 2 1; do $TR --nproc-per-node $n --master-port 2951$n tools/probes/d2h_probe.py --iters 100 >> gpurun_out/r02c_d2h_probe.jsonl 2>> gpurun_out/r02c_d2h_probe.err; echo probe $n rc=$?; done
tail -c 400 gpurun_out/r02c_bench_8gpu.err gpurun_out/r02c_many_player_8gpu.err; tail -3 gpurun_out/r02c_d2h_probe.err
