set -x
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_p.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02_pytest_gpu_p.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
bash tools/r02_scale.sh 1 p
timeout 100 python tools/learner_trace.py 2>&1 | tail -24 > gpurun_out/r02_learner_timeline_p.txt
