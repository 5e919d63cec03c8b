set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python tools/probe_sym.py 1048576 3:0.01:0 3:0.01:4 3:0:0 3:0:0:GRAVSIM_CUTMODE=select 3:0.01:0 2>&1 | grep -v "^+" | tee gpurun_out/r2_probe_final.txt
timeout 600 python bench.py --workload ref65536 > gpurun_out/r2_config1_ref65536.json 2> gpurun_out/r2_config1.err; cut -c1-200 gpurun_out/r2_config1_ref65536.json
