timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -12
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544"
echo "== 2 GPUs default (symmetric + peer exchange)"; timeout 300 $RUN tools/run_scaling_configs.py sweep 65536,262144,1048576 2>&1 | grep -E "^#|Error|error" | head
echo "== 2 GPUs deterministic"; GRAVSIM_FLAGS=128 timeout 300 $RUN tools/run_scaling_configs.py sweep 262144,1048576 2>&1 | grep -E "^#|Error|error" | head
echo "== 1 GPU"; python tools/perf_probe.py 1048576 0 sym 2>&1 | tail -5
