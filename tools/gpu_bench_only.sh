P=$1
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1 --master-port 29544"
timeout 300 $RUN bench.py --gpus $P --steps 5 --warmup 3 > gpurun_out/bench_p$P.json 2> gpurun_out/bench_p$P.err; tail -1 gpurun_out/bench_p$P.err | cut -c1-200
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/bench_p$P.json") if l.startswith("{")][-1])
print("N_GPU=%d value=%.4e ms/step=%.3f pct/gpu=%.2f e2e=%.4e launches=%d variant=%s pass=%s" % (d["n_gpus"], d["value"], d["ms_per_step"], d["pct_fp32_peak"], d["e2e"]["value"], d["gpu_launches"], d["config"]["variant"], d["config"]["force_pass"]))
PY
