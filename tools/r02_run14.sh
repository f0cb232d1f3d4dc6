set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_f.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02_pytest_gpu_f.log
HB_CHAIN_TRACE=1 timeout 600 python -m pytest tests/test_qnet_gpu.py -m gpu -x -q -s -k "learner_step_parity" 2>&1 | grep -E "chain job|expand|c51" | head -10 > gpurun_out/r02_chain_trace_f.txt
timeout 300 python tools/learner_probe.py > gpurun_out/r02_learner_probe_f.txt 2>&1; cat gpurun_out/r02_learner_probe_f.txt
timeout 300 python tools/learner_trace.py 2>&1 | tail -24 > gpurun_out/r02_learner_timeline_f.txt; cat gpurun_out/r02_learner_timeline_f.txt
bash tools/r02_scale.sh 1 f
