# Evidence run on N GPUs of one box (N = $1, default 8): headline bench (with the tree compared to a single-GPU build),
# per-level exchange sweeps (configs 3, 5, 4) and the forest mode against the reference's MPI strategy (config 2).
set -u
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
mkdir -p gpurun_out
timeout 300 $TR --master-port 29601 bench.py --gpus $N --steps 8 --warmup 3 2>gpurun_out/b.err | tail -1 > gpurun_out/bench_r2_2p28_${N}gpu.json
timeout 200 $TR --master-port 29602 tools/exchange_sweep.py 28 3 2>gpurun_out/s1.err | grep -v "^NCCL" > gpurun_out/sweep_r2_2p28_3d_${N}gpu.json
if [ "$N" = "8" ]; then
timeout 200 $TR --master-port 29603 tools/exchange_sweep.py 24 8 2>gpurun_out/s2.err | grep -v "^NCCL" > gpurun_out/sweep_r2_2p24_8d_${N}gpu.json
timeout 200 $TR --master-port 29604 tools/exchange_sweep.py 26 3 clustered 2>gpurun_out/s3.err | grep -v "^NCCL" > gpurun_out/sweep_r2_2p26_3d_clustered_${N}gpu.json
timeout 300 $TR --master-port 29606 tools/config2_gpu.py 24 3 1 2>/dev/null | tail -1 > gpurun_out/config2_2p24_3d_w${N}.json
fi
cut -c1-330 gpurun_out/bench_r2_2p28_${N}gpu.json
