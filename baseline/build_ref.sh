#!/bin/bash
# Builds the UNMODIFIED reference sources against the shim headers.
# baseline/_ref is git-ignored; it holds a verbatim copy of /root/reference (so that the GPU box,
# which has no /root/reference, gets the sources through gpurun) plus the two binaries.
set -e
cd "$(dirname "$0")/.."
mkdir -p baseline/_ref
if [ -d /root/reference ]; then cp /root/reference/kernel.cu /root/reference/kern.cpp baseline/_ref/; fi
test -f baseline/_ref/kernel.cu || { echo "baseline/_ref/kernel.cu missing (no /root/reference here)"; exit 3; }
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
$NVCC -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I baseline/ref_shim -Xcompiler -mssse3 \
    -o baseline/_ref/ref_kernel_cu baseline/_ref/kernel.cu -lrt -lpthread
${CXX:-g++} -O3 -mssse3 -std=c++17 -I baseline/ref_shim -I /usr/local/cuda/include \
    -o baseline/_ref/ref_kern_cpp baseline/_ref/kern.cpp -lrt -lpthread
echo "built baseline/_ref/ref_kernel_cu and baseline/_ref/ref_kern_cpp"
