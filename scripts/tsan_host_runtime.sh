#!/bin/bash
# Builds the host side of the native runtime with ThreadSanitizer (SAN=thread, default) or AddressSanitizer (SAN=address)
# and runs tests/native/tsan_host.cc (no GPU needed). Output is kept in profiles/r2_${SAN}san_host.txt.
set -e
SAN=${SAN:-thread}
cd "$(dirname "$0")/.."
mkdir -p build/${SAN}san
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O1 -g -std=c++17 --expt-relaxed-constexpr -Xcompiler -fsanitize=${SAN} -Xcompiler -fPIC -I include"
for f in collectives elementwise runtime; do
  $NVCC $FLAGS -c csrc/$f.cu -o build/${SAN}san/$f.o
done
g++ -O1 -g -std=c++17 -fsanitize=${SAN} -c tests/native/tsan_host.cc -o build/${SAN}san/tsan_host.o
$NVCC -gencode arch=compute_100a,code=sm_100a -Xcompiler -fsanitize=${SAN} build/${SAN}san/*.o -o build/${SAN}san/tsan_host -lcudart_static -lpthread -ldl -lrt $([ $SAN = thread ] && echo -ltsan || echo -lasan)
TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0" ASAN_OPTIONS="detect_leaks=0:protect_shadow_gap=0" ./build/${SAN}san/tsan_host 2>&1 | tee profiles/r2_$([ $SAN = thread ] && echo t || echo a)san_host.txt
