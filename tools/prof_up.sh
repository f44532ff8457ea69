for m in up down; do
CMD="python tools/time_upfirdn.py $m 32 64 128 f32 6"
timeout 120 $CMD 2>&1 | tail -1 && \
timeout 500 ncu --set full --clock-control none --import-source on -k regex:upfirdn2d_fast -s 6 -c 1 -f -o gpurun_out/prof_up_$m $CMD > gpurun_out/ncu_up_$m.log 2>&1
tail -1 gpurun_out/ncu_up_$m.log
done
python tools/time_upfirdn.py up 32 64 128 bf16 | tail -1
