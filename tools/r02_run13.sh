timeout 900 python -m pytest tests/test_replay_gpu.py tests/test_session_gpu.py tests/test_agent_gpu.py -m gpu -x -q 2>&1 | tail -5
timeout 600 python tools/session_trace.py train 2>&1 | grep -E "tick:|synchronising"
timeout 600 python tools/session_trace.py learn 2>&1 | grep -E "tick:"
