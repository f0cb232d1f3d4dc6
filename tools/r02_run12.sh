timeout 900 python -m pytest tests/test_qnet_gpu.py tests/test_rainbow_gpu.py tests/test_replay_gpu.py tests/test_agent_gpu.py tests/test_session_gpu.py -m gpu -x -q 2>&1 | tail -4
for d in 0 1; do
echo "== dense_head=$d"
HB_DENSE_HEAD_BW=$d timeout 300 python tools/learner_probe.py 2>&1 | grep -E "device events|phase B|hb_learner_step"
HB_DENSE_HEAD_BW=$d timeout 300 python tools/learner_trace.py 2>&1 | tail -30
done
