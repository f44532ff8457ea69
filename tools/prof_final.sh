# round evidence (one gpurun call, one GPU): launch list of the bench command, ncu --set full of the fused filtered_lrelu kernels
# (streaming kernel at the bench shape, whole-plane kernel at 8x8), torch-profiler breakdown of the training step.
# Every ncu run follows a plain run of the same command that exited 0.   usage: bash tools/prof_final.sh [tag]
tag=${1:-r2_final}
set -x
CMD="python bench.py --steps 3 --warmup 3 --no-train"
$CMD > gpurun_out/${tag}_plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_bench.csv $CMD > gpurun_out/${tag}_ncu_launch.log 2>&1
CMD="python tools/time_flr.py 32 64 128 128 f32 6"
TIME_DB=1 timeout 120 $CMD > gpurun_out/${tag}_plain_flr.log 2>&1 && \
TIME_DB=1 timeout 500 ncu --set full --clock-control none --import-source on -k regex:flr_fused -s 12 -c 2 -f -o gpurun_out/${tag}_flr $CMD > gpurun_out/${tag}_ncu_flr.log 2>&1
CMD="python tools/time_flr.py 32 256 8 8 f32 6"
TIME_DB=1 timeout 120 $CMD > gpurun_out/${tag}_plain_plane.log 2>&1 && \
TIME_DB=1 timeout 500 ncu --set full --clock-control none --import-source on -k regex:flr_plane -s 12 -c 2 -f -o gpurun_out/${tag}_plane $CMD > gpurun_out/${tag}_ncu_plane.log 2>&1
timeout 300 python tools/prof_train_torch.py 0 32 > gpurun_out/${tag}_train_step_breakdown.txt 2>&1
tail -n 3 gpurun_out/${tag}_plain_flr.log gpurun_out/${tag}_plain_plane.log
ls -la gpurun_out/*.ncu-rep
