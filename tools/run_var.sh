for v in "$@"; do echo "== $v"; CNO_B200_LIB=$PWD/tools/variants/$v.so timeout 120 python tools/time_flr.py 2>&1 | tail -1; done
