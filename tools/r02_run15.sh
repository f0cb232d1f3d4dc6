set -x
# launch list of the bench (final build) and full captures of the new kernels
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r02_launches_bench_h.csv python bench.py --steps 3 --warmup 3 --no-gate --cpu-seconds 0 --learner-steps 4 --act-steps 2 --tree-iters 20 --e2e-steps 4 --session-ticks 3 > gpurun_out/ncu_launches_h.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"mlp_chain_kernel|c51_loss_kernel|sample_gather_kernel|adam_refresh_kernel" -s 8 -c 6 -o gpurun_out/r02_learner_h -f python tools/learner_probe.py > gpurun_out/ncu_learner_h.log 2>&1; echo "ncu learner rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:selfplay -s 6 -c 1 -o gpurun_out/r02_selfplay_h -f python bench.py --steps 8 --warmup 3 --no-gate --e2e-steps 4 --cpu-seconds 0 --no-learner --no-tree --no-session > gpurun_out/ncu_selfplay_h.log 2>&1; echo "ncu selfplay rc=$?"
ls -la gpurun_out/*.ncu-rep
