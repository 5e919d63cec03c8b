# round 2, evidence on 4 GPUs: multi-GPU tests (2, 3, 4 ranks), bench + N sweep at P = 2 and P = 4
set -x
timeout 1200 python -m pytest tests -m gpu -v -k "multi_gpu or two_gpus" 2>&1 | grep -E "PASSED|FAILED|SKIPPED|passed|failed" | sed 's/^.*:://' | tee gpurun_out/r2_pytest_multi_gpu_4gpu.txt
for n in 2 4; do
  RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533"
  timeout 900 $RUN bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r2_bench_${n}gpu.json 2> gpurun_out/r2_bench_${n}gpu.err; echo "rc=$?"; cut -c1-260 gpurun_out/r2_bench_${n}gpu.json
  timeout 1500 $RUN bench.py --gpus $n --workload sweep > gpurun_out/r2_sweep_p$n.json 2> gpurun_out/r2_sweep_p$n.err; grep "^# sweep" gpurun_out/r2_sweep_p$n.err
done
