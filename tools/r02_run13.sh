timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python tools/learner_probe.py 2>&1 | grep -E "device events"
HB_FWD_CHAIN=0 timeout 300 python tools/learner_probe.py 2>&1 | grep -E "device events"
timeout 600 python tools/session_trace.py train 2>&1 | grep -E "tick:"
timeout 600 python tools/session_trace.py eval 2>&1 | grep -E "tick:"
