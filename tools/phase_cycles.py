#!/usr/bin/env python
"""Where dependency_kernel's cycles go, per phase (development tool; needs lib/librpf_b200_timers.so = the library built
with -DRPF_PHASE_TIMERS, see raytracer-rpf_b200/Makefile). Thread 0 of every CTA reads clock64() at the phase boundaries.

  RPF_B200_LIB=raytracer-rpf_b200/lib/librpf_b200_timers.so python tools/phase_cycles.py [boxes...]
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ.setdefault("RPF_B200_LIB", os.path.join(ROOT, "raytracer-rpf_b200", "lib", "librpf_b200_timers.so"))
sys.path.insert(0, os.path.join(ROOT, "raytracer-rpf_b200"))
import rpf_b200  # noqa: E402

NAMES = ["prepare+decode", "gather+binning", "marginals+clear", "joint fill+sweep", "MI+weights", "-", "(team0 wait after fill)", "(team0 sweep)"]
W, H, P = 640, 360, 8
for box in [int(b) for b in sys.argv[1:]] or [55, 35, 17]:
    planes = rpf_b200.synth_planes(W, H, P, seed=1234, kind=0, x_off=320, y_off=180)
    f = rpf_b200.RPFFilter(W, H, P)
    f.upload(planes)
    f.apply([box], 0.002)
    f.synchronize()
    out = (C.c_ulonglong * 8)()
    L = rpf_b200.lib()
    L.rpf_debug_phase_cycles.argtypes = [C.c_void_p, C.c_void_p]
    L.rpf_debug_phase_cycles(f._h, out)  # warm-up pass discarded
    f.reset_colors()
    f.set_profiling(True)
    f.reset_stats()
    f.apply([box], 0.002)
    f.synchronize()
    kt = f.kernel_times()
    L.rpf_debug_phase_cycles(f._h, out)
    st = f.stats()
    tot = sum(out[i] for i in range(5))
    print(f"box {box}: mean N {st['nbh_sum'] / st['pixels_filtered']:.0f}, dependency_kernel {kt['dependency_kernel'][0]:.1f} ms")
    for i in range(8):
        if out[i]:
            print(f"    {NAMES[i]:28s} {100.0 * out[i] / tot:5.1f} %")
    f.close()
