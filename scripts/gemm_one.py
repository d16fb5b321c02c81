import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from kgp_b200 import _cabi
c = _cabi.Context(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
k = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
print(c.bench_dgemm(n, k, 1), c.bench_dgemm(n, k, 1, syrk=True))
