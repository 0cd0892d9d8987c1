import json, sys
sys.path.insert(0, '/root/repo')
import bench
from ar_orbslam2_b200.matcher import measure_popc_rate
print(json.dumps(bench.bench_c3(None, measure_popc_rate()), indent=1))
