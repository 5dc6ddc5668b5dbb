set -u
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29601 bench.py --gpus 8 --steps 10 --warmup 3 2>gpurun_out/b8.err | tail -1 > gpurun_out/bench_r2_2p28_8gpu.json
timeout 200 $TR --master-port 29602 tools/exchange_sweep.py 28 3 > gpurun_out/sweep_r2_2p28_3d_8gpu.json 2>gpurun_out/s1.err
timeout 200 $TR --master-port 29603 tools/exchange_sweep.py 24 8 > gpurun_out/sweep_r2_2p24_8d_8gpu.json 2>gpurun_out/s2.err
timeout 200 $TR --master-port 29604 tools/exchange_sweep.py 26 3 clustered > gpurun_out/sweep_r2_2p26_3d_clustered_8gpu.json 2>gpurun_out/s3.err
timeout 200 $TR --master-port 29605 bench.py --gpus 8 --steps 5 --warmup 3 --log2n 26 --dist clustered 2>/dev/null | tail -1 > gpurun_out/bench_r2_2p26_clustered_8gpu.json
timeout 300 $TR --master-port 29606 tools/config2_gpu.py 24 3 1 2>/dev/null | tail -1 > gpurun_out/config2_2p24_3d_w8.json
timeout 200 $TR --master-port 29607 tools/dist_check.py 100003x3 4194304x3 300000x8 clustered:2097152x3 2>&1 | grep -E "OK|MISMATCH|backend|ms" > gpurun_out/dist_check_8gpu_symm.log
cut -c1-400 gpurun_out/bench_r2_2p28_8gpu.json; grep -c OK gpurun_out/dist_check_8gpu_symm.log; cat gpurun_out/config2_2p24_3d_w8.json | cut -c1-300
