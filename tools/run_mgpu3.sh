N=$1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
s=$(date +%s); timeout 200 $TR bench.py --gpus $N --steps 20 --warmup 5 2>&1 | grep -E '^\{|Error|error' | cut -c1-200; echo "rc=$? wall=$(( $(date +%s) - s ))s"
s=$(date +%s); timeout 200 $TR bench.py --gpus $N --workload cno_train --steps 10 --warmup 3 --graph 1 2>&1 | grep -E '^\{|Error|error' | cut -c1-200; echo "rc=$? wall=$(( $(date +%s) - s ))s"
s=$(date +%s); timeout 200 $TR bench.py --impl reference --gpus $N --steps 3 --warmup 1 2>&1 | grep -E '^\{|Error|error' | cut -c1-200; echo "wall=$(( $(date +%s) - s ))s"
