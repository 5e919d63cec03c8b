# round 2, call 2 (1 GPU): shuffle transport of the j-side sums vs the shared-memory round trip
set -x
timeout 600 python -m pytest tests -m gpu -x -q -k "symmetric or cutoff or 1M or 65536 or uniform_mass or config1 or one_step" 2>&1 | tail -5 | tee gpurun_out/r2c2_pytest.txt
P="python tools/probe_sym.py 1048576"
timeout 900 $P 3:0.01:0 3:0.01:0:GRAVSIM_SYM_SHFL=0 3:0.01:4 3:0:0 3:0:0:GRAVSIM_CUTMODE=select \
   3:0.01:0:GRAVSIM_SYM_NPAIR=2 3:0.01:4:GRAVSIM_SYM_NPAIR=2 3:0:0:GRAVSIM_SYM_NPAIR=2 \
   17:0.01:0 17:0.01:4 17:0:0 17:0.01:0:GRAVSIM_SYM_SHFL=0 \
   7:0.01:0 7:0.01:0:GRAVSIM_SYM_NPAIR=2 7:0:0 6:0.01:0 6:0.01:0:GRAVSIM_SYM_NPAIR=2 1:0.01:0 2>&1 | grep -v "^+" | tee gpurun_out/r2c2_probe.txt
