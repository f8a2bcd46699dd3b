#!/bin/bash
# Race / memory checking (the reference only has `valgrind memcheck`, unitest/valgrind.sh).
#   tools/sanitize.sh tsan        ThreadSanitizer build of the C++ unit tests (CPU, no GPU needed)
#   tools/sanitize.sh asan        AddressSanitizer build of the C++ unit tests
#   tools/sanitize.sh memcheck    compute-sanitizer --tool memcheck on the single-GPU kernel tests (B200)
#   tools/sanitize.sh racecheck   compute-sanitizer --tool racecheck on the tcgen05 kernel test  (B200)
# (one compute-sanitizer tool per GPU call — see B200_PROFILING.md)
set -e
ROOT="$(cd "$(dirname "$0")/.." && pwd)"; cd "$ROOT"
LZ=$(python -c "from swiftsnails_b200.build import find_libzmq; print(find_libzmq())")
case "${1:-tsan}" in
  tsan|asan)
    F=$([ "$1" = asan ] && echo address || echo thread)
    g++ -O1 -g -std=c++17 -pthread -fsanitize=$F -I include -I csrc/core csrc/tests/unit_tests.cc -o build/unit_tests_$1 \
        -L"$(dirname "$LZ")" -l:"$(basename "$LZ")" -Wl,-rpath,"$(dirname "$LZ")"
    TSAN_OPTIONS="halt_on_error=0 second_deadlock_stack=1" ./build/unit_tests_$1 "${2:-}" ;;
  memcheck)
    compute-sanitizer --tool memcheck --error-exitcode 1 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "not word2vec and not lr_worker" ;;
  racecheck)
    compute-sanitizer --tool racecheck --error-exitcode 1 python -m pytest tests/test_gpu_sgns_tc.py -m gpu -x -q -k "128-300" ;;
esac
