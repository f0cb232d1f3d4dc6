set -x
timeout 600 python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 tools/comm_check.py > gpurun_out/r02_comm_check_n8.log 2>&1; echo "comm rc=$?"
grep -E "ok|all-reduce|learner|Error|assert" gpurun_out/r02_comm_check_n8.log | head
bash tools/r02_scale.sh 8 d
