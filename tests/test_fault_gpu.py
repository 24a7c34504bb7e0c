"""The bounded device-side wait of k_md_steps (a CTA that never takes its ticket must become an error the host raises, not a
hung GPU).  Needs a library built with the fault hook, which the shipped one does not carry:

    make -C hoomd_b200/csrc VARIANT=-DB200_FAULT_INJECT OBJDIR=build_fault OUT=../libhoomd_b200_fault.so
    HOOMD_B200_LIB=$PWD/hoomd_b200/libhoomd_b200_fault.so python -m pytest tests/test_fault_gpu.py -m gpu

With the shipped library the test is skipped."""
import ctypes as C
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_lost_cta_becomes_an_error_and_the_next_system_runs():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import hoomd_b200 as hb
    lib = hb._lib.lib()
    try:
        inject = lib.b200_debug_fault_lose_cta
    except AttributeError:
        pytest.skip("library built without -DB200_FAULT_INJECT")
    inject.argtypes = [C.c_uint]
    inject.restype = C.c_int
    pos, vel, L = hb.workloads.sc_lattice(24, (1.0 / 0.2) ** (1.0 / 3.0), kT=1.2, seed=5)
    pos[:, :3] += np.random.default_rng(6).normal(0, 0.1, (len(pos), 3)).astype(np.float32)

    def make():
        s = hb.System(pos, (L, L, L), vel4=vel)
        s.setup_lj(r_cut=3.0, r_buff=0.4)
        s.setup_integrator(0.005)
        s.add_nve()
        s.set_multistep(True)
        return s

    s = make()
    s.run(40)
    assert s.stepping()["multistep"] is True
    assert inject(3) == 0
    t0 = time.perf_counter()
    with pytest.raises(RuntimeError, match="never reached the grid barrier"):
        s.run(40)
    assert time.perf_counter() - t0 < 30.0     # bounded: 2^22 polls, not a hang
    s.close()
    # a fresh system is unaffected, and runs the multi-step kernel again
    s2 = make()
    s2.run(80)
    assert s2.stepping()["multistep"] is True and np.isfinite(s2.thermo()["pe"])
    s2.close()
