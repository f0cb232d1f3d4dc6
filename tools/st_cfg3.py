#!/usr/bin/env python
"""BASELINE configs[2] alone (capacity 2^22, batch 256 stratified sample + priority update): the
loop of bench.py's bench_sum_tree, for an ncu capture of st_descend_kernel / st_update_kernel."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench

ap = argparse.ArgumentParser()
ap.add_argument("--tree-iters", type=int, default=200)
args = ap.parse_args()
print(json.dumps(bench.bench_sum_tree(args)))
