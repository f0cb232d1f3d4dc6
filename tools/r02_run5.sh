# round 2: full single-GPU evidence of the current build
set -x
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_n1_b_steps20.json 2> gpurun_out/r02_bench_n1_b.err; echo "bench20 rc=$?"
timeout 300 python bench.py --steps 4096 --warmup 16 --no-learner --no-tree --no-session --cpu-seconds 0 > gpurun_out/r02_bench_n1_b_steps4096.json 2>> gpurun_out/r02_bench_n1_b.err; echo "bench4096 rc=$?"
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r02_bench_ref_b.json 2>> gpurun_out/r02_bench_n1_b.err; echo "ref rc=$?"
# launch list of the whole bench (cold caches, serialised)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r02_launches_bench_b.csv python bench.py --steps 3 --warmup 3 --no-gate --cpu-seconds 0 --learner-steps 4 --act-steps 2 --tree-iters 20 --e2e-steps 4 --session-ticks 3 > gpurun_out/ncu_launches_b.log 2>&1; echo "ncu list rc=$?"
# full capture of the headline kernel (one isolated launch)
timeout 600 ncu --set full --clock-control none --import-source on -k regex:selfplay -s 6 -c 1 -o gpurun_out/r02_selfplay_b -f python bench.py --steps 8 --warmup 3 --no-gate --e2e-steps 4 --cpu-seconds 0 --no-learner --no-tree --no-session > gpurun_out/ncu_selfplay_b.log 2>&1; echo "ncu selfplay rc=$?"
# full captures of the GEMM variants: acting pass (3 layers at 65,536 rows) and one learner update
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_f16_kernel -s 9 -c 3 -o gpurun_out/r02_gemm_acting_b -f python bench.py --steps 3 --warmup 3 --no-gate --e2e-steps 4 --cpu-seconds 0 --no-tree --no-session --learner-steps 4 --act-steps 2 > gpurun_out/ncu_gemm_acting_b.log 2>&1; echo "ncu gemm rc=$?"
timeout 600 ncu --set full --clock-control none --cache-control none --import-source on -k regex:"gemm_f16_kernel|sample_gather|c51_loss|adam_refresh" -s 400 -c 16 -o gpurun_out/r02_learner_b -f python tools/learner_probe.py > gpurun_out/ncu_learner_b.log 2>&1; echo "ncu learner rc=$?"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
echo done
