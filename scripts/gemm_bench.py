import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from kgp_b200 import _cabi
c = _cabi.Context(0)
print('dmma peak', c.bench_dmma_peak(20000))
for n, k in [(8192, 8192), (16384, 512), (40064, 512), (40064, 128)]:
    print('gemm n=%d k=%d' % (n, k), c.bench_dgemm(n, k, 3), 'syrk', c.bench_dgemm(n, k, 3, syrk=True), flush=True)
