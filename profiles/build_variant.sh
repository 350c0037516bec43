#!/bin/bash
# profiles/build_variant.sh NAME "-DFLAG=..." [file.cu ...] — in the container: an A/B build of libb200attn.so with extra
# nvcc flags on the named sources (default attn_streamk.cu), the other objects from the product build; written to
# gpurun_variants/NAME/libb200attn.so and selected at run time with B200_LIB=<that path>.
set -e
NAME=$1; FLAGS=$2; shift 2
SRCS=${@:-attn_streamk.cu}
cd "$(dirname "$0")/../llama.cpp_test_b200/csrc"
make -s -j16 all
mkdir -p build/var_$NAME ../../gpurun_variants/$NAME
OBJS=""
for f in shim kv_write attn_generic attn_streamk attn_prefill peer_gather ggml_entry attn_static_ws; do
  if [[ " $SRCS " == *" $f.cu "* ]]; then
    /usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-fvisibility=hidden \
      -I../../include -I. --expt-relaxed-constexpr -Xptxas -v $FLAGS -c $f.cu -o build/var_$NAME/$f.o 2> build/var_$NAME/$f.ptxas.log
    OBJS="$OBJS build/var_$NAME/$f.o"
  else
    OBJS="$OBJS build/$f.o"
  fi
done
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../gpurun_variants/$NAME/libb200attn.so $OBJS build/host_*.o
echo "gpurun_variants/$NAME/libb200attn.so"
