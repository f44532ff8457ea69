# model test + fresh ncu full capture of the fused kernel (fwd WRITE + bwd READ), source pages
python -m pytest tests/test_cno_model.py -m gpu -q --timeout 300 2>&1 | tail -2
CMD="python tools/time_flr.py 32 64 128 128 f32 6"
timeout 120 $CMD > gpurun_out/plain_flr.log 2>&1 && \
timeout 500 ncu --set full --clock-control none --import-source on -k regex:flr_fused -s 12 -c 2 -f -o gpurun_out/prof_flr_cur $CMD > gpurun_out/ncu_full_cur.log 2>&1
tail -2 gpurun_out/ncu_full_cur.log; cat gpurun_out/plain_flr.log | tail -1
