"""One direct measurement of the CPU arm (oracle gp() + dlikExact, all host cores) at a size close to the stated config, to check
the cost model bench.py extrapolates with.  usage: python scripts/cpu_direct.py 32768"""
import sys, os, json, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
bench.use_all_cores()
bench.cpu_eval_seconds(512)
t0 = time.time()
tot, t3 = bench.cpu_eval_seconds(n, split=True)
print(json.dumps({'n': n, 'seconds': tot, 'seconds_blas3': t3, 'cores': bench.cpu_threads(), 'cpu_model': bench.cpu_model(),
                  'wall': time.time() - t0}), flush=True)
