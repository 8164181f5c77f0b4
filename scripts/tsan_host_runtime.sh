#!/bin/bash
# Builds the host side of the native runtime with ThreadSanitizer and runs tests/native/tsan_host.cc (no GPU needed).
# Output is kept in profiles/r2_tsan_host.txt.
set -e
cd "$(dirname "$0")/.."
mkdir -p build/tsan
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O1 -g -std=c++17 --expt-relaxed-constexpr -Xcompiler -fsanitize=thread -Xcompiler -fPIC -I include"
for f in collectives elementwise runtime; do
  $NVCC $FLAGS -c csrc/$f.cu -o build/tsan/$f.o
done
g++ -O1 -g -std=c++17 -fsanitize=thread -c tests/native/tsan_host.cc -o build/tsan/tsan_host.o
$NVCC -gencode arch=compute_100a,code=sm_100a -Xcompiler -fsanitize=thread build/tsan/*.o -o build/tsan/tsan_host -lcudart_static -lpthread -ldl -lrt -ltsan
TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0" ./build/tsan/tsan_host 2>&1 | tee profiles/r2_tsan_host.txt
