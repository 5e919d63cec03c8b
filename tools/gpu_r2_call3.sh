# round 2, call 3 (1 GPU): R = 12/16 builds, real A/B of the knobs, ncu full capture of the shuffle kernel
set -x
timeout 600 python -m pytest tests -m gpu -x -q -k "every_kernel_variant" 2>&1 | tail -3
P="python tools/probe_sym.py 1048576"
timeout 900 $P 3:0.01:0 3:0.01:0:GRAVSIM_SYM_SHFL=0 3:0.01:0:GRAVSIM_SYM_NPAIR=2 3:0.01:0:GRAVSIM_SYM_NPAIR=2,GRAVSIM_SYM_SHFL=0 \
   19:0.01:0 19:0.01:0:GRAVSIM_SYM_SHFL=0 19:0.01:4 19:0:0 18:0.01:0 18:0.01:0:GRAVSIM_SYM_SHFL=0 18:0.01:4 18:0:0 \
   17:0.01:0 17:0.01:0:GRAVSIM_SYM_SHFL=0 7:0.01:0 7:0.01:0:GRAVSIM_SYM_NPAIR=2 3:0.01:4 3:0:0 2>&1 | grep -v "^+" | tee gpurun_out/r2c3_probe.txt
CMD="python tools/probe_sym.py 1048576 3:0.01:0"
ncu --set full --clock-control none --import-source on -k regex:sym_force -s 2 -c 1 -o gpurun_out/r2c3_shfl $CMD > gpurun_out/r2c3_ncu.log 2>&1
echo "ncu rc=$?"
