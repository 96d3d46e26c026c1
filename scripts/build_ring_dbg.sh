#!/bin/bash
# developer tool: build/libdbg.so = the library with the ring kernel's MMA-warp wait-site timing compiled in
set -e
cd "$(dirname "$0")/.."
python -c "import raygun_b200._build as b; b.build()"
O=raygun_b200/lib/obj
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -DRING_DEBUG_TIMING \
  -c raygun_b200/csrc/ring_conv_sm100.cu -o build/ring_dbg.o
nvcc -shared -o build/libdbg.so build/ring_dbg.o $(ls $O/*.o | grep -v ring_conv_sm100) \
  -gencode arch=compute_100a,code=sm_100a -cudart static
