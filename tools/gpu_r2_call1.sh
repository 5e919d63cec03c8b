# round 2, call 1 (1 GPU): GPU tests of the reworked symmetric kernel, then timing probes at N = 1M
set -x
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/r2c1_pytest.txt
P="python tools/probe_sym.py 1048576"
timeout 600 $P 3:0.01:0 3:0.01:4 3:0:0 3:0:0:GRAVSIM_CUTMODE=select \
   3:0.01:0:GRAVSIM_SYM_PROBE=1 3:0.01:0:GRAVSIM_SYM_PROBE=2 3:0.01:0:GRAVSIM_SYM_PROBE=3 \
   17:0.01:0 17:0.01:4 17:0:0 \
   7:0.01:0 7:0.01:0:GRAVSIM_SYM_NPAIR=2 7:0:0 \
   3:0.01:0:GRAVSIM_SYM_NPAIR=2 6:0.01:0 2>&1 | grep -v "^+" | tee gpurun_out/r2c1_probe.txt
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2c1_bench.json 2> gpurun_out/r2c1_bench.err; cut -c1-300 gpurun_out/r2c1_bench.json
