# launch list of the eager CNO_time training step, then the one-step breakdown (tools/launch_breakdown.py).
# BOUNDED on purpose: an unbounded capture (-c 60000 over set-up + 4 steps, inner timeout 600 s) ran into gpurun's limit in
# round 1 and used up the remaining GPU budget.  ~2,500 launches per step: skip the set-up and the first step, keep two steps.
CMD="python bench.py --workload cno_train --steps 1 --warmup 2 --graph 0"
timeout 120 $CMD > gpurun_out/train_plain.log 2>&1 && \
timeout 240 ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 6000 --csv --log-file gpurun_out/launches_train.csv $CMD > gpurun_out/ncu_train.log 2>&1
tail -n 2 gpurun_out/ncu_train.log
python tools/launch_breakdown.py gpurun_out/launches_train.csv > gpurun_out/cno_train_step_breakdown.txt 2>&1; head -60 gpurun_out/cno_train_step_breakdown.txt
