timeout 900 python -m pytest tests/test_agent_gpu.py -m gpu -x -q 2>&1 | tail -8
HB_BWD_CHAIN=0 HB_FWD_CHAIN=0 HB_DENSE_HEAD_BW=1 HB_DZ0_PLAIN=1 timeout 900 python -m pytest tests/test_agent_gpu.py -m gpu -x -q -k against_the_oracle 2>&1 | tail -3
