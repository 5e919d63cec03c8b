set -x
timeout 600 python -m pytest tests -m gpu -x -q -k "symmetric_pass or cutoff or 1M" 2>&1 | tail -3
P="python tools/probe_sym.py 1048576"
timeout 900 $P 3:0.01:0 3:0.01:0:GRAVSIM_SYM_SHFL=0 3:0.01:0:GRAVSIM_SYM_NPAIR=2 3:0.01:0:GRAVSIM_SYM_NPAIR=2,GRAVSIM_SYM_SHFL=0 \
   19:0.01:0 19:0.01:0:GRAVSIM_SYM_SHFL=0 18:0.01:0 17:0.01:0 17:0.01:0:GRAVSIM_SYM_SHFL=0 7:0.01:0 7:0.01:0:GRAVSIM_SYM_NPAIR=2 \
   3:0.01:4 3:0:0 17:0.01:4 17:0:0 19:0.01:4 19:0:0 6:0.01:0 6:0.01:0:GRAVSIM_SYM_NPAIR=2 2>&1 | grep -v "^+" | tee gpurun_out/r2c4_probe.txt
