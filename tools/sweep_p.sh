# planes-per-CTA / persistent-CTA sweep for the mid and small CNO levels (graph timed)
for shape in "32 32 64 64" "32 64 32 32" "32 128 16 16"; do
  for P in 0 4 7 10; do
    for CT in 0 296; do
      echo -n "P=$P CTAS=$CT : "; CNO_FLR_P=$P CNO_FLR_CTAS=$CT timeout 100 python tools/time_flr_graph.py $shape 2>&1 | tail -1
    done
  done
done
