for np in 2 4; do for v in 6 1; do
  echo "== NPAIR=$np variant=$v"; GRAVSIM_SYM_NPAIR=$np python tools/perf_probe.py 1048576 $v sym 2>&1 | grep "SYM"
done; done
GRAVSIM_SYM_NPAIR=4 timeout 300 python -m pytest tests -m gpu -x -q -k "symmetric or 1M" 2>&1 | tail -3
