// Unity translation unit of libkgp_b200.so: the kernels in the .cuh headers are
// shared by the exact path, the grid path and the C ABI, so they are compiled once.
#include "exact.cu"
#include "grid.cu"
#include "api.cu"
