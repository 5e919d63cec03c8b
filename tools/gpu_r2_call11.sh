set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
timeout 300 python tools/probe_sym.py 16384 0:0.01:0 0:0:0 2>&1 | grep -v "^+"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --workload sweep --sweep-max 1048576 > gpurun_out/r2c11_sweep2.json 2> gpurun_out/r2c11_sweep2.err; grep "^# sweep" gpurun_out/r2c11_sweep2.err
timeout 600 python bench.py --workload sweep --sweep-max 1048576 > gpurun_out/r2c11_sweep1.json 2> gpurun_out/r2c11_sweep1.err; grep "^# sweep" gpurun_out/r2c11_sweep1.err
