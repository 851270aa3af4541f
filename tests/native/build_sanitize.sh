#!/bin/sh
# Builds tests/native/sanitize_host twice (ASAN+UBSAN, TSAN) from the runtime's host sources and runs both.
set -e
cd "$(dirname "$0")/../.."
OUT=${1:-/tmp/pdj-sanitize}
mkdir -p "$OUT"
SRC="paddle_operator_b200/csrc/comm_runtime.cu paddle_operator_b200/csrc/tma_util.cu tests/native/sanitize_host.cpp"
for tag in address thread; do
  if [ $tag = address ]; then FLAGS="-Xcompiler -fsanitize=address -Xcompiler -fsanitize=undefined"; LINK="-Xlinker -lasan -Xlinker -lubsan";
  else FLAGS="-Xcompiler -fsanitize=thread"; LINK="-Xlinker -ltsan"; fi
  nvcc -gencode arch=compute_100a,code=sm_100a -O1 -g -std=c++17 -I include -I paddle_operator_b200/csrc \
       $FLAGS -Xcompiler -fno-omit-frame-pointer $LINK -cudart shared \
       $SRC -o "$OUT/sanitize_host_$tag" -lpthread
  # CUDA maps large fixed address ranges: ASAN must not reserve the shadow gap; leak checking of the driver is noise
  ASAN_OPTIONS=protect_shadow_gap=0:detect_leaks=0 TSAN_OPTIONS=report_signal_unsafe=0 "$OUT/sanitize_host_$tag"
done
echo "sanitize: ok"
