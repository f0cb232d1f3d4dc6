for r in 1 2; do for v in 1 0; do echo "HB_MERGED_TRANSPOSE=$v"; HB_MERGED_TRANSPOSE=$v timeout 300 python tools/learner_probe.py 2>&1 | grep -E "device events"; done; done
