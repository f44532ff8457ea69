# one ncu --set full capture (with SASS source counters) of the fused filtered_lrelu forward + backward launch at the bench
# shape, after the same command exited 0 without ncu.  usage: tools/prof_flr.sh <tag> [N C S_in S_out dtype]
tag=${1:-flr}; shift
ARGS=${@:-32 64 128 128 f32}
CMD="python tools/time_flr.py $ARGS 6"
TIME_DB=1 timeout 120 $CMD > gpurun_out/plain_$tag.log 2>&1 && \
TIME_DB=1 timeout 500 ncu --set full --clock-control none --import-source on -k regex:flr_fused -s 12 -c 2 -f -o gpurun_out/$tag $CMD > gpurun_out/ncu_$tag.log 2>&1
tail -n 3 gpurun_out/plain_$tag.log gpurun_out/ncu_$tag.log
