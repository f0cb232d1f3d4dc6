timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_qnet_gpu.py tests/test_rainbow_gpu.py tests/test_agent_gpu.py tests/test_session_gpu.py -m gpu -x -q 2>&1 | tail -2
HB_CHAIN_TRACE=1 timeout 600 python -m pytest tests/test_qnet_gpu.py -m gpu -x -q -s -k "learner_step_parity and 658-hidden0" 2>&1 | grep -E "chain job [45]" | tail -2 | cut -c1-200
for i in 1 2; do timeout 300 python tools/learner_probe.py 2>&1 | grep -E "device events"; done
timeout 600 python bench.py --steps 3 --warmup 3 --e2e-steps 4 --cpu-seconds 0 --no-tree --no-session --learner-steps 200 --act-steps 20 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('acting', d['acting']['rows_per_s'], d['acting']['ms_per_call'], d['acting']['roofline']['frac'])
print('learner', d['learner']['updates_per_s'])"
