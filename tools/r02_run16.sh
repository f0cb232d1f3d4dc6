set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_n.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02_pytest_gpu_n.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
bash tools/r02_scale.sh 1 n
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r02_bench_ref_n.json 2> gpurun_out/r02_bench_ref_n.err; echo "ref rc=$?"
HB_LOSS_TRACE=1 HB_CHAIN_TRACE=1 timeout 600 python -m pytest tests/test_qnet_gpu.py -m gpu -x -q -s -k "learner_step_parity" 2>&1 | grep -E "chain job|expand|c51" | head -12 > gpurun_out/r02_chain_loss_trace_n.txt
timeout 300 python tools/learner_trace.py 2>&1 | tail -24 > gpurun_out/r02_learner_timeline_n.txt
timeout 300 python tools/session_trace.py train 2>&1 | grep -v Warn | tail -34 > gpurun_out/r02_session_timeline_n.txt
timeout 300 python tools/learner_probe.py > gpurun_out/r02_learner_probe_n.txt 2>&1; cat gpurun_out/r02_learner_probe_n.txt
