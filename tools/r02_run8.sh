set -x
timeout 900 python -m pytest tests/test_qnet_gpu.py tests/test_rainbow_gpu.py tests/test_replay_gpu.py tests/test_agent_gpu.py tests/test_session_gpu.py -m gpu -x -q > gpurun_out/r02_gputest_g.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/r02_gputest_g.log
timeout 300 python tools/learner_probe.py > gpurun_out/r02_learner_probe_d.log 2>&1; echo "probe rc=$?"
cat gpurun_out/r02_learner_probe_d.log
HB_DENSE_HEAD_BW=1 timeout 300 python tools/learner_probe.py > gpurun_out/r02_learner_probe_d_dense.log 2>&1; echo "probe rc=$?"
cat gpurun_out/r02_learner_probe_d_dense.log
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"gemm_f16_kernel.*256" -c 3 -o gpurun_out/r02_gemm_acting_e -f python bench.py --steps 3 --warmup 3 --no-gate --e2e-steps 4 --cpu-seconds 0 --no-tree --no-session --learner-steps 4 --act-steps 2 > gpurun_out/r02_gemm_acting_e.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/r02_gemm_acting_e.log | cut -c1-300
