# usage: bash tools/gpu_scaling.sh P "N,N,..." [galaxyN galaxySteps]   (run under gpurun --gpus P)
P=$1; NS=$2
if [ "$P" -eq 1 ]; then RUN="python"; else RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $P --master-addr 127.0.0.1 --master-port 29544"; fi
$RUN tools/run_scaling_configs.py sweep $NS > gpurun_out/sweep_p$P.json 2> gpurun_out/sweep_p$P.err; grep "^#" gpurun_out/sweep_p$P.err
if [ -n "$3" ]; then
  $RUN tools/run_scaling_configs.py galaxy $3 $4 > gpurun_out/galaxy_p$P.json 2> gpurun_out/galaxy_p$P.err; tail -c 1500 gpurun_out/galaxy_p$P.json
fi
