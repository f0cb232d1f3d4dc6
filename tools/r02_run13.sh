for r in 1 2; do for v in 1 0; do echo "HB_ADAM0_PDL=$v"; HB_ADAM0_PDL=$v timeout 300 python tools/learner_probe.py 2>&1 | grep -E "device events"; done; done
timeout 600 python -m pytest tests/test_agent_gpu.py tests/test_qnet_gpu.py -m gpu -x -q 2>&1 | tail -2
