for shape in "32 256 8 8" "32 64 32 32"; do
CMD="python tools/time_flr.py $shape f32 6"
timeout 120 $CMD 2>&1 | tail -1
tag=$(echo $shape | tr ' ' '_')
timeout 500 ncu --set full --clock-control none --import-source on -k regex:flr_fused -s 12 -c 2 -f -o gpurun_out/prof_flr_$tag $CMD > gpurun_out/ncu_small_$tag.log 2>&1
tail -1 gpurun_out/ncu_small_$tag.log
done
