set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
P="python tools/probe_sym.py 1048576"
timeout 900 $P 3:0.01:0 3:0.01:4 3:0:0 2>&1 | grep -v "^+" | tee gpurun_out/r2c8_probe.txt
timeout 300 python tools/probe_sym.py 65536 3:0:0 3:0.01:0 2>&1 | grep -v "^+"
