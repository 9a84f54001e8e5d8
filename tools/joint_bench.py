"""Timing of the joint N-vehicle RSU edge solve alone (bench.measure_joint): python tools/joint_bench.py"""
import json, os, sys, types
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R)
import bench
r = bench.measure_joint(types.SimpleNamespace(steps=20), 0)
print(json.dumps({k: v for k, v in r.items() if k.startswith("S")}))
