set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputest_e.log 2>&1; echo "pytest rc=$?"
tail -12 gpurun_out/r02_gputest_e.log
echo done
