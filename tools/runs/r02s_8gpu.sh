# 8-GPU lease with the final round-2 build: the driver's own bench command at 8 ranks and BASELINE config 5 at 8 x 8 192 worlds.
# Run as: gpurun --gpus 8 -- 'bash tools/runs/r02s_8gpu.sh'  (outputs under gpurun_out/)
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
mkdir -p gpurun_out
$TR --nproc-per-node 8 --master-port 29501 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r02s_bench_8gpu.json 2> gpurun_out/r02s_bench_8gpu.err; echo bench8 rc=$?
$TR --nproc-per-node 8 --master-port 29502 bench.py --gpus 8 --steps 20 --warmup 5 --workload many_player --no-cpu-baseline > gpurun_out/r02s_many_player_8gpu.json 2> gpurun_out/r02s_many_player_8gpu.err; echo many_player8 rc=$?
grep -h '^{' gpurun_out/r02s_bench_8gpu.json gpurun_out/r02s_many_player_8gpu.json | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['config']['workload'][:40], d['n_gpus'], d['value'], d['roofline']['frac'], d.get("e2e", {}).get("value"))"
