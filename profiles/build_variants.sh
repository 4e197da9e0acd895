#!/bin/bash
# Builds libdcv variants with different warp splits of the fused tail kernel into deconv_api_b200/_lib/variants/
# (A/B runs: DCV_LIB_PATH=<file> python profiles/ab.py ...).  usage: build_variants.sh "16 4" "8 8" ...
set -e
cd "$(dirname "$0")/../deconv_api_b200/csrc"
mkdir -p ../_lib/variants ../_lib/obj
for cfg in "$@"; do
  set -- $cfg
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden \
       -DSTRIP_EPI_WARPS_CFG=$1 -DSTRIP_PROD_WARPS_CFG=$2 -I ../../include -c conv_tc.cu -o ../_lib/variants/conv_tc_$1_$2.o
  nvcc -shared -o ../_lib/variants/libdcv_$1_$2.so ../_lib/obj/api.o ../_lib/obj/kernels_stream.o ../_lib/obj/conv_fp32.o \
       ../_lib/variants/conv_tc_$1_$2.o ../_lib/obj/noalloc.o ../_lib/obj/dream.o ../_lib/obj/dev_ops.o -gencode arch=compute_100a,code=sm_100a
  echo built libdcv_$1_$2.so
done
