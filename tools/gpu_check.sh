set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r2_pytest_gpu_1gpu.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_1gpu.json 2> gpurun_out/r2_bench_1gpu.err; tail -2 gpurun_out/r2_bench_1gpu.err; cut -c1-300 gpurun_out/r2_bench_1gpu.json
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-cuda-baseline --no-checks"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "ncu1 rc=$?"
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sym_force -s 3 -c 1 -f -o gpurun_out/prof $CMD > gpurun_out/ncu2.log 2>&1
echo "ncu2 rc=$?"
