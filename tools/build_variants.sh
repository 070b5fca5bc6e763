#!/bin/bash
# A/B builds of the text kernels: libtmog_<tag>.so under gpurun_out/variants (development aid)
set -e
cd "$(dirname "$0")/../transmogrifai_b200/csrc"
mkdir -p ../../variants build
build() {  # tag, extra nvcc flags
  tag=$1; shift
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-O3 --expt-relaxed-constexpr "$@" -c text.cu -o build/text_$tag.o
  objs=$(ls build/*.o | grep -v "build/text" | tr '\n' ' ')
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../variants/libtmog_$tag.so $objs build/text_$tag.o -cudart static -ldl
}
for v in "$@"; do
  case $v in
    rl16) build rl16 -DTC_ROUND_LANES_V=16 ;;
    rl20) build rl20 -DTC_ROUND_LANES_V=20 ;;
    rl30) build rl30 -DTC_ROUND_LANES_V=30 ;;
    p6) build p6 -DTC_PEND_V=6 ;;
    p20) build p20 -DTC_PEND_V=20 ;;
    w4) build w4 -DTC_WARPS_V=4 ;;
    w12) build w12 -DTC_WARPS_V=12 ;;
    w16) build w16 -DTC_WARPS_V=16 ;;
    ir8k) build ir8k -DTS_ITEM_ROWS_V=8192 ;;
    ir4k) build ir4k -DTS_ITEM_ROWS_V=4096 ;;
    ir1k) build ir1k -DTS_ITEM_ROWS_V=1024 ;;
    sw4) build sw4 -DTS_WARPS_V=4 ;;
    sw6) build sw6 -DTS_WARPS_V=6 ;;
    sw12) build sw12 -DTS_WARPS_V=12 ;;
    sw24) build sw24 -DTS_WARPS_V=24 ;;
  esac
done
