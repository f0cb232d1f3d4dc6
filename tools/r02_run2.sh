set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputest_d.log 2>&1; echo "pytest rc=$?"
tail -8 gpurun_out/r02_gputest_d.log
timeout 300 python tools/learner_probe.py > gpurun_out/r02_learner_probe_b.log 2>&1; echo "probe rc=$?"
cat gpurun_out/r02_learner_probe_b.log
HB_LEARNER_PDL=0 timeout 300 python tools/learner_probe.py > gpurun_out/r02_learner_probe_b_nopdl.log 2>&1; echo "probe rc=$?"
cat gpurun_out/r02_learner_probe_b_nopdl.log
echo done
