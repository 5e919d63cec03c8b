timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
echo "== tracked (new default)"; python tools/perf_probe.py 1048576 3,6,7,1 cut 2>&1 | grep "mode=cut"
echo "== exact select"; GRAVSIM_EXACT_SELECT=1 python tools/perf_probe.py 1048576 3,6 cut 2>&1 | grep "mode=cut"
echo "== tracked, N=65536"; python tools/perf_probe.py 65536 3,6 cut 2>&1 | grep "mode=cut"
