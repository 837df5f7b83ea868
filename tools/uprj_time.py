"""comp-uprjdists builder leg alone (k_desc-bound): 100k pairs x 68 filters; prints pairs/s and the kernel split."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dlco_b200, bench_paths
ctx = dlco_b200.Context(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
print(json.dumps(bench_paths.run_uprj_builder(ctx, n, 68, bench_paths._peaks()[0])))
