CMD="python tools/time_flr.py 32 64 128 128 f32 6"
timeout 120 $CMD > gpurun_out/plain_flr.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on -k regex:flr_fused -s 12 -c 2 -o gpurun_out/prof_flr_r1c $CMD > gpurun_out/ncu_full3.log 2>&1
tail -2 gpurun_out/ncu_full3.log
