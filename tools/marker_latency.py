#!/usr/bin/env python3
"""Marker-path front end alone (bench.py: extra.marker_orb): cv::ORB latency per frame + ratio matcher, CPU legs beside."""
import json
import os
import sys
import types

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

if __name__ == "__main__":
    print(json.dumps(bench.bench_marker(types.SimpleNamespace(skip_cpu="--skip-cpu" in sys.argv)), indent=1))
