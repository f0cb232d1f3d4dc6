for r in 1 2; do for v in 1 0; do echo "HB_ACT_T_SIDE=$v"; HB_ACT_T_SIDE=$v timeout 300 python tools/learner_probe.py 2>&1 | grep -E "device events"; done; done
timeout 200 python -m pytest tests/test_agent_gpu.py tests/test_qnet_gpu.py -x -q -m gpu 2>&1 | tail -3
