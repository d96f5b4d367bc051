"""mpmc_b200_recip_energy_static (include/mpmc_b200.h): the reciprocal-space Ewald energy for callers that cannot
allocate -- scratch in a file-scope device array, no cudaMalloc inside the call.  No counterpart in the reference.
Called through the C ABI (ctypes) with caller-owned device buffers and held to
mpmc_b200_energy()'s es_recip (itself pinned to coulombic_reciprocal(), src/energy/coulombic.c:42-95, by the parity
tests) at the suite's tolerance, on calls in a row that share the array: large system, small one (stale contents of the
large one must not leak), the large one again.  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

from mpmc_b200 import systems
from mpmc_b200.api import load_library

pytestmark = pytest.mark.gpu

RTOL = 1e-10  # tests/golden_util.py: north_star's tolerance for each Ewald piece


def _static_recip(lib, torch, s, stream=0, check_free=True):
    p = s.params
    dev = torch.device("cuda", 0)
    d_pos = torch.from_numpy(s.pos.reshape(-1)).to(dev)
    d_q = torch.from_numpy(s.charge).to(dev)
    d_frozen = torch.from_numpy(s.frozen).to(dev)
    d_out = torch.full((1,), float("nan"), dtype=torch.float64, device=dev)
    rb = (C.c_double * 9)(*[p.reciprocal_basis[i][j] for i in range(3) for j in range(3)])
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(0)[0]
    rc = lib.mpmc_b200_recip_energy_static(
        C.c_int(s.n), C.c_void_p(d_pos.data_ptr()), C.c_void_p(d_q.data_ptr()), C.c_void_p(d_frozen.data_ptr()), rb,
        C.c_double(p.volume), C.c_double(p.ewald_alpha), C.c_int(p.ewald_kmax), C.c_void_p(d_out.data_ptr()),
        C.c_void_p(stream))
    assert rc == 0, lib.mpmc_b200_last_error().decode()
    torch.cuda.synchronize()
    if check_free:
        assert torch.cuda.mem_get_info(0)[0] == free_before, "the call changed the amount of free device memory"
    return float(d_out.cpu()[0])


def test_static_scratch_entry_point_matches_energy(engine):
    import torch

    torch.cuda.set_device(0)
    lib = load_library()
    big = systems.co2_fluid(4096)  # 12,288 charged mobile atoms: more sub-chunks than CTAs
    small = systems.mixed_small(triclinic=True)  # triclinic cell, frozen charged atoms, uncharged atoms
    big7 = systems.mof_h2(cells=(3, 3, 1))  # frozen framework + H2
    want = {}
    for s in (big, small, big7):
        engine.set_system(s)
        want[s.name] = engine.energy().es_recip
        assert want[s.name] != 0.0
    torch.cuda.set_device(0)
    # the first call on a device is the set-up call: the driver loads the module's kernels and instantiates the static
    # array then (lazy loading), which shows as a one-time drop of free memory; no later call may change it
    _static_recip(lib, torch, small, check_free=False)
    for s in (big, small, big7, big, small):  # calls in a row on the shared array
        got = _static_recip(lib, torch, s)
        assert abs(got - want[s.name]) <= RTOL * abs(want[s.name]), (s.name, got, want[s.name])
    # on a caller's stream
    st = torch.cuda.Stream()
    got = _static_recip(lib, torch, big, stream=st.cuda_stream)
    assert abs(got - want[big.name]) <= RTOL * abs(want[big.name])


def test_static_scratch_entry_point_rejects_sizes_beyond_its_limits():
    import torch

    torch.cuda.set_device(0)
    lib = load_library()
    one = torch.zeros(8, dtype=torch.float64, device="cuda:0")
    rb = (C.c_double * 9)(*([0.1, 0, 0, 0, 0.1, 0, 0, 0, 0.1]))

    def call(n, kmax):
        return lib.mpmc_b200_recip_energy_static(C.c_int(n), C.c_void_p(one.data_ptr()), C.c_void_p(one.data_ptr()), None,
                                                 rb, C.c_double(1000.0), C.c_double(0.7), C.c_int(kmax),
                                                 C.c_void_p(one.data_ptr()), None)

    assert call((1 << 20) + 1, 7) == 3
    assert b"1048576" in lib.mpmc_b200_last_error()
    assert call(2, 8) == 3
    assert b"ewald_kmax" in lib.mpmc_b200_last_error()
    assert call(2, 7) == 0
    torch.cuda.synchronize()
