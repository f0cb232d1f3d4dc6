set -x
timeout 900 python -m pytest tests/test_replay_gpu.py tests/test_agent_gpu.py tests/test_session_gpu.py -m gpu -x -q > gpurun_out/r02_gputest_f.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/r02_gputest_f.log
timeout 300 python tools/learner_probe.py > gpurun_out/r02_learner_probe_c.log 2>&1; echo "probe rc=$?"
cat gpurun_out/r02_learner_probe_c.log
# the acting pass's GEMMs (BN = 256 tiles are only used at 65,536 rows): full capture
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"gemm_f16_kernel<., 256" -c 3 -o gpurun_out/r02_gemm_acting_c -f python bench.py --steps 3 --warmup 3 --no-gate --e2e-steps 4 --cpu-seconds 0 --no-tree --no-session --learner-steps 4 --act-steps 2 > gpurun_out/ncu_gemm_acting_c.log 2>&1; echo "ncu gemm rc=$?"
echo done
