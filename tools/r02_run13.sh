timeout 900 python -m pytest tests/test_session_gpu.py tests/test_agent_gpu.py -m gpu -x -q 2>&1 | tail -3
timeout 600 python tools/session_trace.py learn 2>&1 | grep -E "tick:"
timeout 600 python tools/session_trace.py train 2>&1 | grep -E "tick:"
