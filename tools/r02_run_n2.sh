set -x
timeout 600 python -m pytest tests/test_comm_gpu.py -m gpu -x -q > gpurun_out/r02_comm_test.log 2>&1; echo "comm rc=$?"
tail -4 gpurun_out/r02_comm_test.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02_bench_n2_a.json 2> gpurun_out/r02_bench_n2_a.err; echo "bench2 rc=$?"
tail -3 gpurun_out/r02_bench_n2_a.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02_bench_n2_a.json'))
print('value %.4g'%d['value'], 'frac', d['roofline']['frac'], 'e2e %.4g'%d['e2e']['value'])
print({k:v for k,v in d['learner'].items() if k!='cpu_baseline'})
print(d['session'])
PY
echo done
