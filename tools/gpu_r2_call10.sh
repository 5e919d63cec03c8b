set -x
for n in 4096 8192 16384 32768 65536 131072; do
 timeout 300 python tools/probe_sym.py $n 10:0.01:0 10:0.01:128 16:0.01:128 1:0.01:0 3:0.01:0 0:0.01:0 10:0:0 10:0:128 3:0:0 2>&1 | grep -v "^+"
done | tee gpurun_out/r2c10_smallN.txt
