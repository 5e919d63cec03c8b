# 8 GPUs: multi-GPU oracle tests for 2..8 ranks (odd and even), the 1M-body all-GPU check, bench at 8
set -x
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 1500 python -m pytest tests -m gpu -q -k "multi_gpu or two_gpus" -v 2>&1 | grep -E "PASS|FAIL|SKIP|passed|failed|Error" | tee gpurun_out/r2_pytest_multi_gpu_8.txt
n=8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/r2c7_bench_n$n.json 2> gpurun_out/r2c7_bench_n$n.err
echo "rc=$? n=$n"; tail -2 gpurun_out/r2c7_bench_n$n.err; python -c "
import json; d=json.loads([l for l in open('gpurun_out/r2c7_bench_n$n.json') if l.startswith('{')][-1]); print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['checks'])"
