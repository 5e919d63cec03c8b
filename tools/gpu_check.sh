timeout 300 python tools/probe_sym.py 1048576 3:0.01:0 3:0.01:4 3:0:0 3:0.01:0 2>&1 | grep -v "^+"
timeout 300 python -m pytest tests -m gpu -x -q -k "symmetric or cutoff or reproducible" 2>&1 | tail -3
