N=$1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 200 $TR bench.py --gpus $N --steps 20 --warmup 5 2>&1 | grep -E '^\{|Error|error' | cut -c1-900
timeout 300 $TR bench.py --gpus $N --workload cno_train --steps 10 --warmup 3 --graph 1 2>&1 | grep -E '^\{|Error|error' | cut -c1-700
