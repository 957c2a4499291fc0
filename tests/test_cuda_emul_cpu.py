"""Checks of the CUDA emulation itself (tests/host_emul/cuda_emul): the warp / block primitives against closed forms
(including threads that exit before a barrier), and the schedule probe — a kernel with a missing __syncthreads gives
schedule-dependent results under GSR_EMUL_SCHED=reverse, the same kernel with the barrier does not."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
EMU = os.path.join(HERE, "host_emul", "cuda_emul")
OUT = os.path.join(HERE, "host_emul", "libemul_selftest.so")


def build():
    srcs = [os.path.join(EMU, f) for f in ("selftest.cu", "emul.cpp", "cuda_runtime.h")]
    if not os.path.exists(OUT) or os.path.getmtime(OUT) < max(os.path.getmtime(f) for f in srcs):
        subprocess.check_call(["g++", "-std=c++17", "-O1", "-shared", "-fPIC", "-U_FORTIFY_SOURCE", "-D_FORTIFY_SOURCE=0",
                               "-DGSR_HOST_EMUL", "-I" + EMU, "-x", "c++", srcs[0], "-x", "c++", srcs[1], "-o", OUT])
    return OUT


@pytest.mark.parametrize("n,block", ((256, 256), (200, 128), (77, 64), (1, 32)))
def test_primitives(n, block):
    lib = C.CDLL(build())
    out = np.zeros(8 * n + 1, np.uint32)
    lib.run_primitives(out.ctypes.data_as(C.c_void_p), C.c_int(n), C.c_int(block))
    assert out[-1] == n
    o = out[:-1].reshape(n, 8)
    gid = np.arange(n)
    for g in range(n):
        w0 = g - g % 32
        lanes = np.arange(w0, min(w0 + 32, n))                    # live lanes of this warp
        b0 = g - g % block
        blk = np.arange(b0, min(b0 + block, n))
        full = len(lanes) == 32
        if full:                                                  # xor-butterfly sums are defined for full warps
            assert o[g, 0] == lanes.sum()
        assert o[g, 1] == (g if g == w0 else g - 1)
        assert o[g, 2] == sum(1 << (l - w0) for l in lanes if l & 1)
        assert o[g, 3] == w0
        assert o[g, 4] == sum(1 << (l - w0) for l in lanes if l % 3 == g % 3)
        assert o[g, 5] == (blk & 1).sum()
        assert o[g, 7] == (1 if 5 in lanes else 0) | 2
    assert gid.size == n


SCRIPT = """
import ctypes as C, numpy as np, sys
lib = C.CDLL(sys.argv[1])
a = np.arange(256, dtype=np.uint32) + 1000
for barrier in (1, 0):
    out = np.zeros(256, np.uint32)
    lib.run_reverse(a.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), C.c_int(barrier))
    print(int(np.array_equal(out, a[::-1])))
"""


def test_schedule_probe_exposes_a_missing_barrier():
    so = build()
    res = {}
    for mode in ("", "reverse", "random:3"):
        env = dict(os.environ, GSR_EMUL_SCHED=mode)
        r = subprocess.run([sys.executable, "-c", SCRIPT, so], env=env, capture_output=True, text=True, check=True)
        res[mode] = [int(x) for x in r.stdout.split()]
    assert all(v[0] == 1 for v in res.values())                   # with the barrier: right under every schedule
    # without it a thread reads slots their owners have not written yet: the emulation must show the bug
    assert any(v[1] == 0 for v in res.values())
