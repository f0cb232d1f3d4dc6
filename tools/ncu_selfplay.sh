set -x
timeout 300 python bench.py --steps 64 --warmup 32 --e2e-steps 4 --cpu-seconds 0 --no-learner --no-tree --no-session > gpurun_out/plain_sp.log 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:selfplay -s 2 -c 1 -o gpurun_out/r01_selfplay_v4 -f python bench.py --steps 64 --warmup 32 --e2e-steps 4 --cpu-seconds 0 --no-learner --no-tree --no-session > gpurun_out/ncu_sp.log 2>&1
echo done
