set -x
timeout 900 python -m pytest tests/test_sum_tree_gpu.py tests/test_replay_gpu.py tests/test_agent_gpu.py -m gpu -x -q > gpurun_out/r02_gputest_tree.log 2>&1; echo "pytest rc=$?"
tail -6 gpurun_out/r02_gputest_tree.log
timeout 200 python tools/st_cfg3.py --tree-iters 2000 > gpurun_out/r02_st_cfg3_b.log 2>&1; echo "st rc=$?"
cat gpurun_out/r02_st_cfg3_b.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"st_descend|st_update_kernel" -s 40 -c 4 -o gpurun_out/r02_st_cfg3 -f python tools/st_cfg3.py --tree-iters 60 > gpurun_out/ncu_st_cfg3.log 2>&1; echo "ncu st rc=$?"
echo done
