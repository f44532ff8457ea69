# ncu evidence for the CIN -> filtered_lrelu fused pair: statistics kernel, AFF forward variant, the backward kernels
CMD="python tools/time_pair.py 32 64 128 f32 6"
timeout 120 $CMD > gpurun_out/plain_pair.log 2>&1 && \
timeout 500 ncu --set full --clock-control none --import-source on -k regex:'cin_|flr_fused|film_' -s 16 -c 5 -f -o gpurun_out/r1_pair $CMD > gpurun_out/ncu_pair.log 2>&1
tail -n 2 gpurun_out/plain_pair.log; tail -n 2 gpurun_out/ncu_pair.log
python tools/ncu_raw_summary.py gpurun_out/r1_pair.ncu-rep > gpurun_out/r1_pair_ncu_full_summary.txt 2>&1; head -50 gpurun_out/r1_pair_ncu_full_summary.txt
