# usage: bash tools/r02_scale.sh N TAG   (under gpurun --gpus N)
N=$1; TAG=$2
set -x
if [ "$N" = "1" ]; then
  timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_n1_$TAG.json 2> gpurun_out/r02_bench_n1_$TAG.err
else
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r02_bench_n${N}_$TAG.json 2> gpurun_out/r02_bench_n${N}_$TAG.err
fi
echo "bench rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/r02_bench_n${N}_$TAG.json'))
print('N', d['n_gpus'], 'value %.4g'%d['value'], 'frac %.3f'%d['roofline']['frac'], 'e2e %.4g'%d['e2e']['value'], 'ceiling', d['e2e'].get('ceiling',{}).get('value'), d['e2e'].get('ceiling',{}).get('copy_GBps_all_ranks'))
l=d['learner']; print('learner', l['updates_per_s'], l['ms_per_update'], l.get('allreduce_impl'), l.get('allreduce_check'), l.get('allreduce_us'), l.get('nccl_allreduce_us'))
print('session', {k:round(v,4) for k,v in d['session'].items() if k.endswith('_ms')})
print('acting', d['acting']['rows_per_s'])
PY
