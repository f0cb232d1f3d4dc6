timeout 600 python -m pytest tests/test_gemm_gpu.py tests/test_qnet_gpu.py tests/test_rainbow_gpu.py tests/test_agent_gpu.py tests/test_session_gpu.py -m gpu -x -q 2>&1 | tail -3
for v in 0 1; do echo "HB_DZ0_PLAIN=$v"; HB_DZ0_PLAIN=$v timeout 300 python tools/learner_probe.py 2>&1 | grep -E "device events|hb_learner_step"; done
