# full GPU suite, N=1 bench, reference arm, acting-pass GEMM capture
set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_e.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02_pytest_gpu_e.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
bash tools/r02_scale.sh 1 e
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r02_bench_ref_e.json 2> gpurun_out/r02_bench_ref_e.err; echo "ref rc=$?"; cat gpurun_out/r02_bench_ref_e.json | cut -c1-400
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"gemm_f16_kernel<., 256" -c 3 -o gpurun_out/r02_gemm_acting_e -f python bench.py --steps 3 --warmup 3 --no-gate --e2e-steps 4 --cpu-seconds 0 --no-tree --no-session --learner-steps 4 --act-steps 2 > gpurun_out/r02_gemm_acting_e.log 2>&1; echo "ncu rc=$?"; tail -5 gpurun_out/r02_gemm_acting_e.log
