# usage: bash tools/gpu_confirm.sh P "N,N,..."   -- bench.py at P GPUs + sweep at P GPUs
P=$1
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1 --master-port 29544"
timeout 600 $RUN bench.py --gpus $P --steps 5 --warmup 3 > gpurun_out/bench_p$P.json 2> gpurun_out/bench_p$P.err; tail -2 gpurun_out/bench_p$P.err | cut -c1-300
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/bench_p$P.json") if l.startswith("{")][-1])
print("N_GPU=%d value=%.4e ms/step=%.3f pct/gpu=%.2f e2e=%.4e launches=%d exchange=%s" % (d["n_gpus"], d["value"], d["ms_per_step"], d["pct_fp32_peak"], d["e2e"]["value"], d["gpu_launches"], d["config"].get("exchange")))
PY
timeout 600 $RUN tools/run_scaling_configs.py sweep $2 > gpurun_out/sweep_p$P.json 2> gpurun_out/sweep_p$P.err; grep "^#" gpurun_out/sweep_p$P.err
