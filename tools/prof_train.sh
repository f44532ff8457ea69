CMD="python bench.py --workload cno_train --steps 1 --warmup 3 --graph 0"
$CMD > gpurun_out/train_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 11600 -c 9000 --csv --log-file gpurun_out/launches_train_r1.csv $CMD > gpurun_out/ncu_train.log 2>&1
tail -2 gpurun_out/ncu_train.log
python -m pytest tests/test_cno_model.py -m gpu -q --timeout 200 2>&1 | grep -E "AssertionError|passed|failed" | cut -c1-400
