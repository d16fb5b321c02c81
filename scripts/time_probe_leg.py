"""Timing of the Chebyshev-Hutchinson probe leg (bench.py kissgp_probe_leg) on one GPU, sorted vs unsorted recurrence."""
import sys, os, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
r = bench.kissgp_probe_leg(int(sys.argv[1]) if len(sys.argv) > 1 else 1000000, 0, 3, 2, 1)
print(json.dumps({k: r[k] for k in ('ms_per_eval', 'logdet_and_derivatives_ms', 'nlZ')}))
