timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
python tools/perf_probe.py 1048576 6 2>&1 | tail -2
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544"
echo "== p2p default"; timeout 300 $RUN tools/run_scaling_configs.py sweep 16384,65536,262144,1048576 2>&1 | grep -E "^#|Error|error" | head
echo "== p2p overlap"; GRAVSIM_FLAGS=8 timeout 300 $RUN tools/run_scaling_configs.py sweep 16384,65536,262144 2>&1 | grep -E "^#|Error|error" | head
echo "== nccl"; GRAVSIM_FLAGS=16 timeout 300 $RUN tools/run_scaling_configs.py sweep 16384,65536,262144 2>&1 | grep -E "^#|Error|error" | head
