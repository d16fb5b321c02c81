#!/bin/bash
# round-2 job 2: GEMM lab (issue-rate variants, TMA kernel vs cp.async kernel), parity tests on the new kernel, n=40k timings
set -x
O=gpurun_out/job2; mkdir -p $O
timeout 120 lab/gemm_lab quick > $O/lab_quick.log 2>&1; echo "rc=$?" >> $O/lab_quick.log
if grep -q "mismatching elements: [1-9]" $O/lab_quick.log || ! grep -q "rc=0" $O/lab_quick.log; then echo "TMA kernel wrong or hung" >> $O/lab_quick.log; export KGP_GEMM_IMPL=0; fi
timeout 600 lab/gemm_lab > $O/lab_full.log 2>&1; echo "rc=$?" >> $O/lab_full.log
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
for impl in 1 0; do
  KGP_GEMM_IMPL=$impl timeout 300 python scripts/time_exact_classes.py 40000 > $O/classes_impl$impl.log 2>&1
done
KGP_GEMM_IMPL=1 KGP_LOOKAHEAD=1 timeout 300 python scripts/time_exact.py 40000 > $O/time_exact_impl1_la1.log 2>&1
# which kernel does cuBLAS run for FP64 8192^3 (name, grid, block, registers, shared memory)?
cat > /tmp/cublas_d.py <<'PY'
import torch
a = torch.randn(8192, 8192, dtype=torch.float64, device='cuda'); b = torch.randn(8192, 8192, dtype=torch.float64, device='cuda')
for _ in range(3): c = a @ b.T
torch.cuda.synchronize()
PY
python /tmp/cublas_d.py && ncu --set full --clock-control none -k "regex:gemm|cutlass|xmma|Kernel" -s 1 -c 1 -o $O/cublas_dgemm -f python /tmp/cublas_d.py > $O/ncu_cublas.log 2>&1
ncu -i $O/cublas_dgemm.ncu-rep --page raw --csv > $O/cublas_dgemm_raw.csv 2>/dev/null
ls -la $O
