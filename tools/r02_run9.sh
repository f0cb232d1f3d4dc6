set -x
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -s 2500 -c 200 --csv --log-file gpurun_out/r02_launches_learner_sparse.csv python tools/learner_probe.py > gpurun_out/ncu_learner_sparse.log 2>&1; echo "ncu rc=$?"
HB_DENSE_HEAD_BW=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -s 2500 -c 200 --csv --log-file gpurun_out/r02_launches_learner_dense.csv python tools/learner_probe.py > gpurun_out/ncu_learner_dense.log 2>&1; echo "ncu rc=$?"
