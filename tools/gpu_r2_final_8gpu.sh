# round 2, evidence on 8 GPUs: multi-GPU tests for 2..8 ranks, bench and N sweep at P = 8, configs[3] (two galaxies, 500 steps)
set -x
timeout 1500 python -m pytest tests -m gpu -v -k "multi_gpu or two_gpus" 2>&1 | grep -E "PASSED|FAILED|SKIPPED|passed|failed" | sed 's/^.*:://' | tee gpurun_out/r2_pytest_multi_gpu_8gpu.txt
n=8
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533"
timeout 900 $RUN bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r2_bench_${n}gpu.json 2> gpurun_out/r2_bench_${n}gpu.err; echo "rc=$?"; cut -c1-260 gpurun_out/r2_bench_${n}gpu.json
timeout 1500 $RUN bench.py --gpus $n --workload sweep > gpurun_out/r2_sweep_p$n.json 2> gpurun_out/r2_sweep_p$n.err; grep "^# sweep" gpurun_out/r2_sweep_p$n.err
timeout 1500 $RUN bench.py --gpus $n --workload galaxies4m > gpurun_out/r2_config3_two_galaxy_4M_8gpu.json 2> gpurun_out/r2_config3.err; echo "rc=$?"; tail -3 gpurun_out/r2_config3.err; cut -c1-400 gpurun_out/r2_config3_two_galaxy_4M_8gpu.json
