#!/usr/bin/env bash
# AddressSanitizer + UndefinedBehaviorSanitizer over the host-side native code (csrc/host/*.cpp): the BPE trainer / encoder
# and the host AdamW. compute-sanitizer is closed on the GPU pool, so this is the sanitizer evidence the repository can
# produce; it needs no GPU. Writes profiles/sanitize_host.log.
# ThreadSanitizer is not part of it: the parallel loops are OpenMP `parallel for` over disjoint elements, and TSan cannot
# see libgomp's implicit barrier, so it reports the sequential code after each loop as racing with the loop (tried; 46 such
# reports, all of that shape).
set -euo pipefail
cd "$(dirname "$0")/.."
out=${TMPDIR:-/tmp}/b200_sanitize_host
g++ -std=c++17 -O1 -g -fno-omit-frame-pointer -fopenmp -fsanitize=address,undefined -fno-sanitize-recover=all \
    csrc/host/bpe.cpp csrc/host/cpu_adam.cpp scripts/sanitize_host.cpp -o "$out"
{
  echo "# $(g++ --version | head -1); flags: -fsanitize=address,undefined -fno-sanitize-recover=all -fopenmp"
  for t in 1 4; do
    echo "## OMP_NUM_THREADS=$t"
    OMP_NUM_THREADS=$t ASAN_OPTIONS=detect_leaks=1:abort_on_error=1 UBSAN_OPTIONS=print_stacktrace=1 "$out"
  done
} 2>&1 | tee profiles/sanitize_host.log
