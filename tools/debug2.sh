for d in 1 2 3 4 5 0; do echo "--- dbg=$d"; CNO_FLR_DEBUG=$d python tools/debug_gpu.py flr_none 2>&1 | cut -c1-250; done
echo "--- noffma"; CNO_B200_LIB=$PWD/tools/libcno_noffma.so python tools/debug_gpu.py flr_none flr_write bias_act 2>&1 | cut -c1-400
