set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
P="python tools/probe_sym.py 1048576"
timeout 900 $P 3:0.01:0 3:0.01:0:GRAVSIM_SYM_SHFL=0 17:0.01:0 17:0.01:0:GRAVSIM_SYM_SHFL=0 3:0.01:4 3:0:0 17:0.01:4 17:0:0 2>&1 | grep -v "^+" | tee gpurun_out/r2c5_probe.txt
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r2c5_bench.json 2> gpurun_out/r2c5_bench.err; tail -3 gpurun_out/r2c5_bench.err; cat gpurun_out/r2c5_bench.json
timeout 600 python bench.py --workload ref65536 > gpurun_out/r2c5_ref65536.json 2> gpurun_out/r2c5_ref65536.err; tail -3 gpurun_out/r2c5_ref65536.err; cat gpurun_out/r2c5_ref65536.json
