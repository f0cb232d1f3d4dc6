for b in 0 1; do for c in 8 32; do
echo "== HB_BIAS0_MAIN=$b CUDA_DEVICE_MAX_CONNECTIONS=$c"
CUDA_DEVICE_MAX_CONNECTIONS=$c HB_BIAS0_MAIN=$b timeout 300 python tools/learner_probe.py 2>&1 | grep -E "device events|hb_learner_step"
done; done
HB_BIAS0_MAIN=1 timeout 300 python tools/learner_trace.py 2>&1 | tail -14
