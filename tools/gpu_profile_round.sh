# One-GPU evidence run: GPU tests, smoke, bench (with CPU baseline), ncu launch list + full capture of the
# force kernel (each ncu run directly after the same command exited 0 without ncu), microbenchmark, reference configs.
set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/pytest_gpu.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/bench_p1.json 2> gpurun_out/bench_p1.err; tail -2 gpurun_out/bench_p1.err; cut -c1-400 gpurun_out/bench_p1.json
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2>&1; cut -c1-300 gpurun_out/bench_ref.json
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "ncu1 rc=$?"
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sym_force -s 3 -c 1 -o gpurun_out/prof $CMD > gpurun_out/ncu2.log 2>&1
echo "ncu2 rc=$?"
timeout 600 python tests/reference_configs.py > gpurun_out/reference_configs.json 2> gpurun_out/refcfg.err
timeout 300 python tools/perf_probe.py 1048576 0 sym > gpurun_out/probe_1m.txt 2>&1; cat gpurun_out/probe_1m.txt
