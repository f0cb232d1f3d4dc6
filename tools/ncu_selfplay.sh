# usage: tools/ncu_selfplay.sh <tag>   (under gpurun; plain run first, then ONE ncu capture of selfplay_kernel)
TAG=${1:-r01_selfplay}
set -x
timeout 300 python bench.py --steps 64 --warmup 32 --e2e-steps 4 --cpu-seconds 0 --no-learner --no-tree --no-session > gpurun_out/plain_$TAG.log 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:selfplay -s 2 -c 1 -o gpurun_out/$TAG -f python bench.py --steps 64 --warmup 32 --e2e-steps 4 --cpu-seconds 0 --no-learner --no-tree --no-session > gpurun_out/ncu_$TAG.log 2>&1
cp hanabi_agents_b200/_lib/libhanabi_b200.so gpurun_out/$TAG.lib.so
echo done
