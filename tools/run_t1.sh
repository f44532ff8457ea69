for c in 0 148 296 444 592 683 1024; do CNO_FLR_CTAS=$c python tools/time_flr.py 2>&1 | tail -1; done
python tools/time_flr.py 32 16 128 128 | tail -1
python tools/time_flr.py 32 32 64 64 | tail -1
python tools/time_flr.py 32 64 32 32 | tail -1
python tools/time_flr.py 32 256 8 8 | tail -1
python tools/time_flr.py 32 32 128 64 | tail -1
python tools/time_flr.py 32 16 64 128 | tail -1
python tools/time_flr.py 32 64 128 128 bf16 | tail -1
