# 4 GPUs: multi-GPU oracle tests (2, 3, 4 ranks), bench at 2 and 4 ranks, sanitizer runs, small-N probes
set -x
nvidia-smi --query-gpu=index,name --format=csv,noheader
timeout 1200 python -m pytest tests -m gpu -x -q -k "multi_gpu or two_gpus" 2>&1 | tail -8 | tee gpurun_out/r2c6_pytest_multi.txt
for n in 2 4; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/r2c6_bench_n$n.json 2> gpurun_out/r2c6_bench_n$n.err
  echo "rc=$? n=$n"; tail -2 gpurun_out/r2c6_bench_n$n.err; cut -c1-260 gpurun_out/r2c6_bench_n$n.json; python -c "
import json; d=json.loads([l for l in open('gpurun_out/r2c6_bench_n$n.json') if l.startswith('{')][-1]); print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['checks'])"
done



