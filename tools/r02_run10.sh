for d in 0 1; do for p in 0 1; do
echo "== dense_head=$d adam_pdl=$p"
HB_DENSE_HEAD_BW=$d HB_ADAM_PDL=$p timeout 300 python tools/learner_probe.py 2>&1 | grep -E "device events|phase B|hb_learner_step"
done; done
