CMD="python tools/time_upfirdn.py down 32 64 128 bf16 6"
timeout 120 $CMD 2>&1 | tail -1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:upfirdn2d_fast -s 6 -c 1 -f -o gpurun_out/r1_up_down_bf16 $CMD > gpurun_out/ncu_up_bf16.log 2>&1
tail -1 gpurun_out/ncu_up_bf16.log
