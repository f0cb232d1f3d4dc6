set -x
timeout 600 python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/comm_check.py > gpurun_out/r02_comm_check_n2_g.log 2>&1; echo "comm rc=$?"
tail -8 gpurun_out/r02_comm_check_n2_g.log
timeout 600 python -m pytest tests/test_comm_gpu.py -m gpu -x -q 2>&1 | tail -3
bash tools/r02_scale.sh 2 g
