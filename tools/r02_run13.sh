for i in 1 2 3; do timeout 600 python bench.py --steps 20 --warmup 5 --e2e-steps 16 --cpu-seconds 0 --no-tree --no-session --no-learner 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('value %.4g'%d['value'], 'frac %.3f'%d['roofline']['frac'], 'iso ms %.4f'%d['roofline']['isolated_launch_ms'])"; done
timeout 600 python bench.py --steps 4096 --warmup 16 --e2e-steps 16 --cpu-seconds 0 --no-tree --no-session --no-learner 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('4096: value %.4g'%d['value'], 'frac %.3f'%d['roofline']['frac'])"
