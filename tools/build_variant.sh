#!/bin/bash
# tools/build_variant.sh <name> <nvcc defines...>  ->  gpupathtracer_b200/variants/libb200pt_<name>.so (experimental builds for tools/ab_variants.sh)
name=$1; shift
cd "$(dirname "$0")/../gpupathtracer_b200/csrc" && mkdir -p ../variants && nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -prec-div=true -prec-sqrt=true -ftz=false -Xcompiler -fPIC,-O2,-ffp-contract=off,-fno-fast-math,-pthread -shared -ldl "$@" -o ../variants/libb200pt_$name.so pt_api.cu pt_comm.cu pt_noalloc.cu pt_kernels.cu pt_bvh_gpu.cu pt_host.cpp pt_bvh_build.cpp pt_bvh_wide.cpp pt_bvh_opt.cpp pt_obj.cpp pt_image.cpp
