timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python tools/learner_probe.py 2>&1 | grep -E "device events"
