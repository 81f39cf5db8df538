#!/bin/bash
# ThreadSanitizer (and AddressSanitizer with "asan") build of the C++ runtime + the stress harness; runs on the CPU box.
#   dev/tsan_runtime.sh [tsan|asan]   -> profiles/sanitizer_runtime_<kind>_r2.txt
cd "$(dirname "${BASH_SOURCE[0]}")/.."
kind=${1:-tsan}
flag=$([ "$kind" = asan ] && echo "-fsanitize=address,undefined" || echo "-fsanitize=thread")
out=/tmp/rt_stress_$kind
g++ -std=c++17 -O1 -g $flag -fno-omit-frame-pointer -pthread csrc/runtime/arena.cpp csrc/runtime/broker.cpp tests/native/rt_stress.cpp -o $out || exit 1
{ echo "# $kind run of tests/native/rt_stress.cpp over csrc/runtime/*.cpp ($(g++ --version | head -1)), $(date -u +%Y-%m-%dT%H:%MZ)"; TSAN_OPTIONS="halt_on_error=0" $out 2>&1; echo "exit code $?"; } | tee profiles/sanitizer_runtime_${kind}_r2.txt | tail -15
