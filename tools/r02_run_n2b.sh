set -x
timeout 600 python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/comm_check.py > gpurun_out/r02_comm_check_n2.log 2>&1; echo "comm rc=$?"
tail -12 gpurun_out/r02_comm_check_n2.log
echo done
