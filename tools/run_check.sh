# full GPU check: parity suite, bench line, per-op sweep
python -m pytest tests -m gpu -x -q --timeout 600 2>&1 | tail -5 > gpurun_out/pytest_gpu.log
cat gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
python bench.py --sweep > gpurun_out/sweep.log 2>&1; tail -3 gpurun_out/sweep.log
cp profiles/sweep_latest.json gpurun_out/ 2>/dev/null
