set -x
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r02_noalloc_alone.log 2>&1; echo "alone rc=$?"
tail -4 gpurun_out/r02_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputest_c.log 2>&1; echo "all rc=$?"
tail -6 gpurun_out/r02_gputest_c.log
echo done
