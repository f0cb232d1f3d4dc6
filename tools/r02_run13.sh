timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_qnet_gpu.py tests/test_agent_gpu.py -m gpu -x -q 2>&1 | tail -2
timeout 600 python tools/session_trace.py eval 2>&1 | grep -v Warn | tail -5
timeout 600 python bench.py --steps 3 --warmup 3 --e2e-steps 4 --cpu-seconds 0 --no-tree --no-session --learner-steps 50 --act-steps 20 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('acting', d['acting']['rows_per_s'], d['acting']['ms_per_call'], d['acting']['roofline']['frac'])
print('learner', d['learner']['updates_per_s'])"
