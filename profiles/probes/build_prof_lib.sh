#!/bin/sh
# Builds profiles/probes/libtcb200_prof.so: the library with the phase timers compiled in
# (-DTCB_NMS_PROFILE: clock64 phase counters of yolo_nms_kernel, read by profiles/prof_nms.py;
#  -DTCB_SWEEP_PROFILE: per-phase cycle printout of wide::warp_mask_sweep, used by profiles/probes/prof_sweep.py).
# The profiling build is never the one that is benchmarked or tested; it is not committed (*.so is git-ignored).
set -e
cd "$(dirname "$0")/../../triton_client_b200/csrc"
make -j8 > /dev/null
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC,-Wall,-Wno-unused-function -cudart static"
$NV -DTCB_NMS_PROFILE -c yolo_post.cu -o /tmp/tcb_yolo_post_prof.o
$NV -DTCB_SWEEP_PROFILE -c post3d.cu -o /tmp/tcb_post3d_prof.o
OBJS=$(ls *.o | grep -v -e '^yolo_post.o$' -e '^post3d.o$')
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o ../../profiles/probes/libtcb200_prof.so $OBJS /tmp/tcb_yolo_post_prof.o /tmp/tcb_post3d_prof.o
echo "built profiles/probes/libtcb200_prof.so"
