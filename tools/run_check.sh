# full GPU check: parity suite, bench line (+ reference arm), per-op sweep, CNO train line
python -m pytest tests -m gpu -q --timeout 600 2>&1 | tail -3 > gpurun_out/pytest_gpu.log; cat gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; cut -c1-600 gpurun_out/bench.json; tail -3 gpurun_out/bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err; cut -c1-300 gpurun_out/bench_ref.json
python bench.py --sweep > gpurun_out/sweep.log 2>&1; tail -2 gpurun_out/sweep.log
cp profiles/sweep_latest.json gpurun_out/ 2>/dev/null
python bench.py --workload cno_train --steps 10 --warmup 3 > gpurun_out/bench_train.json 2>> gpurun_out/bench.err; cut -c1-400 gpurun_out/bench_train.json
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
CNO_FUSE_LRELU=0 python bench.py --workload cno_train --steps 10 --warmup 3 > gpurun_out/bench_train_unfused.json 2>> gpurun_out/bench.err; cut -c1-200 gpurun_out/bench_train_unfused.json
