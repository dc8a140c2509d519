"""Run tools/call_latency (the C++ call-latency harness) on the synthetic scenes of bench.py and print its JSON.
   python tools/run_call_latency.py            (environment switches such as SLAMGPU_BF_ZEROCOPY=0 pass through)"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import racing_slam_b200 as pkg  # noqa: E402

print(json.dumps(bench.c_abi_call_latency(pkg, keep_inputs="--keep" in sys.argv)))
