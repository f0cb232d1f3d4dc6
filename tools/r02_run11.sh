for d in 0 1; do
echo "== dense_head=$d"
HB_DENSE_HEAD_BW=$d timeout 300 python tools/learner_trace.py 2>&1 | tail -45
done
