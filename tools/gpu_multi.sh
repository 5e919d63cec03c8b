# usage: bash tools/gpu_multi.sh NGPU  -- 2-GPU test + bench at 1..NGPU (run under gpurun --gpus NGPU)
NG=${1:-2}
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 600 python -m pytest tests -m gpu -x -q -k "two_gpus or two_pass" 2>&1 | tail -5
for n in 1 2 4 8; do
  if [ $n -le $NG ]; then
    if [ $n -eq 1 ]; then
      timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
    else
      timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
    fi
    echo "rc=$? n=$n"; tail -2 gpurun_out/bench_n$n.err; python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/bench_n$n.json") if l.startswith("{")][-1])
    print("N_GPU=%d value=%.4e int/s  ms/step=%.3f  pct_peak/gpu=%.2f  e2e=%.4e launches=%d clocks=%s" % (d["n_gpus"], d["value"], d["ms_per_step"], d["pct_fp32_peak"], d["e2e"]["value"], d["gpu_launches"], d["clocks"]))
except Exception as e: print("parse failed", e)
PY
  fi
done
