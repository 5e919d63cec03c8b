# round 2, final evidence on ONE GPU: tests, smoke, bench (+ reference arm), ncu launch list and full capture (each
# after the same command exited 0 without ncu), configs[1], the N sweep at P = 1
set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/r2_pytest_gpu_1gpu.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2_bench_reference_arm.json 2>&1; cut -c1-200 gpurun_out/r2_bench_reference_arm.json
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_1gpu.json 2> gpurun_out/r2_bench_1gpu.err; tail -2 gpurun_out/r2_bench_1gpu.err; cut -c1-300 gpurun_out/r2_bench_1gpu.json
timeout 600 python bench.py --workload ref65536 > gpurun_out/r2_config1_ref65536.json 2> gpurun_out/r2_config1.err; cut -c1-300 gpurun_out/r2_config1_ref65536.json
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-cuda-baseline --no-checks"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "ncu1 rc=$?"
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sym_force -s 3 -c 1 -f -o gpurun_out/prof $CMD > gpurun_out/ncu2.log 2>&1
echo "ncu2 rc=$?"
timeout 900 python bench.py --workload sweep > gpurun_out/r2_sweep_p1.json 2> gpurun_out/r2_sweep_p1.err; grep "^# sweep" gpurun_out/r2_sweep_p1.err
