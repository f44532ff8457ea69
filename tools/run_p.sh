for shape in "32 16 128 128" "32 32 64 64" "32 64 32 32" "32 128 16 16" "32 256 8 8" "32 64 64 32" "32 32 32 64" "32 128 8 16" "32 256 16 8"; do
  timeout 100 python tools/time_flr_graph.py $shape | tail -1
done
