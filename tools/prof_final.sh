# round-1 evidence: launch list of the bench command + one full-set capture per hot kernel (each after a plain run exited 0)
set -x
CMD="python bench.py --steps 3 --warmup 3"
$CMD > gpurun_out/plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_launches_bench.csv $CMD > gpurun_out/ncu_launch.log 2>&1
CMD="python tools/time_flr.py 32 64 128 128 f32 6"
TIME_DB=1 timeout 120 $CMD > gpurun_out/plain_flr.log 2>&1 && \
TIME_DB=1 timeout 500 ncu --set full --clock-control none --import-source on -k regex:flr_fused -s 12 -c 2 -f -o gpurun_out/r1_flr_final $CMD > gpurun_out/ncu_flr_final.log 2>&1
for m in up down; do
CMD="python tools/time_upfirdn.py $m 32 64 128 f32 6"
timeout 120 $CMD > gpurun_out/plain_up_$m.log 2>&1 && \
timeout 500 ncu --set full --clock-control none --import-source on -k regex:upfirdn2d_fast -s 6 -c 1 -f -o gpurun_out/r1_up_$m $CMD > gpurun_out/ncu_up_$m.log 2>&1
done
CMD="python tools/time_cin.py 32 64 128 f32 6"
timeout 120 $CMD > gpurun_out/plain_cin.log 2>&1 && \
timeout 500 ncu --set full --clock-control none --import-source on -k regex:cin_ -s 6 -c 2 -f -o gpurun_out/r1_cin $CMD > gpurun_out/ncu_cin.log 2>&1
CMD="python tools/time_cin.py 8 64 256 f32 6"
timeout 120 $CMD > gpurun_out/plain_cin256.log 2>&1 && \
timeout 500 ncu --set full --clock-control none --import-source on -k regex:cin_ -s 6 -c 2 -f -o gpurun_out/r1_cin256 $CMD > gpurun_out/ncu_cin256.log 2>&1
ls -la gpurun_out/*.ncu-rep
