set -x
timeout 300 python tools/gemm_bench.py 65536 512 512 0 > gpurun_out/plain_g3.log 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_f16_kernel -c 3 -o gpurun_out/r01_gemm_v3 -f python tools/gemm_bench.py 65536 512 512 0 > gpurun_out/ncu_g3.log 2>&1
timeout 300 python tools/bench_session.py 65536 4 0 > gpurun_out/plain_s3.log 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:st_range_internal_scan_kernel -s 30 -c 2 -o gpurun_out/r01_st_scan -f python tools/bench_session.py 65536 4 0 > gpurun_out/ncu_s3.log 2>&1
tail -2 gpurun_out/plain_g3.log gpurun_out/ncu_g3.log gpurun_out/ncu_s3.log
