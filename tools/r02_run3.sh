set -x
# warm-cache per-kernel durations of the learner update (ncu serialises launches; --cache-control none keeps L2 warm)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -s 2500 -c 260 --csv --log-file gpurun_out/r02_launches_learner_warm.csv python tools/learner_probe.py > gpurun_out/ncu_learner_warm.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/ncu_learner_warm.log
echo done
