#!/bin/bash
# Out-of-bounds check of the product's kernels (not part of the pytest run): builds the emulated library
# (all csrc/*.cu against tests/host_emul/cuda_emul) with AddressSanitizer and runs the emulation test files on it.
# Every global buffer the tests pass is a heap allocation (red zones from the preloaded ASan runtime) and every
# __shared__ array a static object (global red zones), so reads or writes past a buffer that a GPU would silently
# tolerate abort here.  SAN=undefined runs UndefinedBehaviorSanitizer instead (shift counts, signed overflow, misaligned
# or null accesses in the kernel sources).  Round 1: 66 tests, no report from either.
# Usage: [SAN=address|undefined] bash tests/host_emul/asan_check.sh
set -e
ROOT=$(cd "$(dirname "$0")/../.." && pwd)
SAN=${SAN:-address}
OUT=${TMPDIR:-/tmp}/gsr_san_$SAN; mkdir -p "$OUT"; cd "$OUT"
EMU=$ROOT/tests/host_emul/cuda_emul; CSRC=$ROOT/gaussian_splatting_b200/csrc
FLAGS="-std=c++17 -O1 -g -fPIC -ffp-contract=off -U_FORTIFY_SOURCE -D_FORTIFY_SOURCE=0 -fsanitize=$SAN -fno-sanitize=vptr -fno-sanitize-recover=all -fno-omit-frame-pointer -DGSR_HOST_EMUL -I$EMU"
for u in api preprocess binning render_fwd render_bwd preprocess_bwd adam ssim exchange densify; do
    g++ $FLAGS -c -x c++ $CSRC/$u.cu -o $u.o &
done; wait
g++ $FLAGS -shared $EMU/emul.cpp *.o -o libgsrast_asan.so
cat > run_pytest.py <<PY
import sys, os; sys.path.insert(0, "$ROOT"); os.chdir("$ROOT")
import ctypes as C
orig = C.CDLL
def cdll(path, *a, **k):
    if isinstance(path, str) and path.endswith(("libgsrast_full_emul.so", "libgsrast_emul.so")): path = "$OUT/libgsrast_asan.so"
    return orig(path, *a, **k)
C.CDLL = cdll
import pytest
sys.exit(pytest.main(["-x", "-q", "-p", "no:cacheprovider", "tests/test_kernels_emul_cpu.py", "tests/test_densify_emul_cpu.py", "tests/test_plyio_cpu.py"]))
PY
if [ "$SAN" = address ]; then
    LD_PRELOAD=$(gcc -print-file-name=libasan.so) ASAN_OPTIONS=detect_leaks=0:detect_stack_use_after_return=0:verify_asan_link_order=0 python run_pytest.py
else
    UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=1 python run_pytest.py
fi
