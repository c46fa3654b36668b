# usage: bash scripts/build_variant.sh <name> [-DFLAG ...]  — build/var/librtc_cuda_<name>.so from the working tree
name=$1; shift
mkdir -p build/var
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo --fmad=false -std=c++17 \
  -Xcompiler -fPIC,-ffp-contract=off,-fvisibility=hidden -shared "$@" \
  -o build/var/librtc_cuda_$name.so raytracer-challenge-rs_b200/csrc/rtc_api.cu
