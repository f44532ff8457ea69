P=tools/probes/tma_probe
run() { echo -n "$* -> "; timeout 60 $P "$@" 2>&1 | tail -1; }
for c in -8 -4 -3 -1 1 3 4 5 8 120 126; do run 128 128 8  32 6 4  $c -6 0; done
