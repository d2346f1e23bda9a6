#!/bin/bash
# Build an A/B variant of libb200attn.so: the persistent prefill kernel recompiled with extra -D flags, all other
# objects from the regular build.  Usage: tools/build_variant.sh NAME "-DPS_W1_LOW=72 -DPS_W1_HIGH=216"
# Result: nano-vllm-v1_b200/b200attn/variants/libb200attn_NAME.so (select with B200ATTN_LIB=<path>).
set -e
NAME=$1; FLAGS=$2; SRC=$3   # optional 3rd argument: another source file for the prefill kernel (e.g. an older revision)
ROOT=$(cd "$(dirname "$0")/.." && pwd)
C=$ROOT/nano-vllm-v1_b200/csrc
mkdir -p $ROOT/nano-vllm-v1_b200/b200attn/variants /tmp/b200_variants
ARCH="-gencode arch=compute_100a,code=sm_100a"
(cd $C && make -s -j8 >/dev/null)
/usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -Xcompiler -fPIC $ARCH --expt-relaxed-constexpr $FLAGS -I$C -I$ROOT/include -c ${SRC:-$C/attn_prefill_persist.cu} -o /tmp/b200_variants/pp_$NAME.o
OBJS=$(ls $C/*.o | grep -v attn_prefill_persist.o)
/usr/local/cuda/bin/nvcc -shared $ARCH -o $ROOT/nano-vllm-v1_b200/b200attn/variants/libb200attn_$NAME.so $OBJS /tmp/b200_variants/pp_$NAME.o -lcudart
echo built $NAME
