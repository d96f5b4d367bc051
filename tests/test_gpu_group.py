"""One process driving several GPUs (mpmc_b200_create_group, the single-process form of the multi-GPU path that sits
behind the drop-in boundary): needs >= 2 CUDA devices, skipped on a 1-GPU box.  The sharded evaluation -- every exchange
done by the engine's own kernels over peer memory, no NCCL, no Python in the loop -- must equal the reference fixtures,
the pins of the benchmarked configurations and the single-GPU evaluation."""
import numpy as np
import pytest

import bench
from mpmc_b200 import Engine, systems
from tests import pins
from tests import golden_util as gu

pytestmark = pytest.mark.gpu


def _ngpu():
    import torch

    return torch.cuda.device_count()


@pytest.fixture(scope="module")
def group():
    n = _ngpu()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 2 if n < 4 else 4
    g = Engine(devices=list(range(world)))
    assert g.group_size() == world
    yield g
    g.close()


FIXTURES = ["mixed_default", "mixed_tri_default", "mixed_palmo", "mixed_sor", "mixed_esor", "mixed_precision", "mixed_rrms",
            "mixed_zodid", "mixed_polar_ewald", "mixed_tri_polar_ewald", "mixed_wolf_a013", "mixed_wolf_full",
            "mixed_damp_linear", "mixed_nopolar_kmax5", "mixed_rd_only", "mixed_es_wolf", "mixed_rd_crystal2",
            "h2_fluid_500_precision_palmo_rrms", "co2_fluid_375", "ref_socmof", "ref_pcn61_h2_500"]


@pytest.mark.parametrize("name", FIXTURES)
def test_group_matches_reference_fixture(group, name):
    s, exp = gu.load(name)
    group.set_system(s)
    got = group.energy()
    gu.assert_energies(got.to_dict(), exp, rtol=1e-10, label=name)
    if s.params.polarization and not s.params.rd_only:
        gu.assert_atom_arrays(group.atom_arrays(), exp, s, rtol=1e-10, label=name)
    again = group.energy()
    assert again.total == got.total and again.polar == got.polar  # rank-ordered sums: reproducible bit for bit


@pytest.mark.parametrize("name", ["pin_mof_h2_21k", "pin_h2_fluid_16000", "pin_co2_fluid_12k", "pin_mof_h2_63k"])
def test_group_matches_pin(group, name):
    if name not in pins.names():
        pytest.skip(f"{name} not generated")
    pin = pins.load(name)
    s = bench.build_workload(pin["meta"]["workload"])
    assert pins.check_inputs(pin, s) == []
    group.set_system(s)
    got = group.energy()
    polar = bool(s.params.polarization and not s.params.rd_only)
    dev = pins.deviation(pin, got.to_dict(), group.atom_arrays() if polar else None)
    assert dev["iterations_equal"] and dev["max"] <= 1e-10, dev
    c = group.counters()
    if polar:
        assert c.pairs_dipole == float(c.n_polarizable) * (c.n_polarizable - 1)


def test_group_follows_moves_and_resizing(group):
    """set_positions reaches every rank; a different atom count re-wires the exchange buffers (uVT insert / remove)."""
    one = Engine(0)
    try:
        s = systems.h2_fluid(3000)
        group.set_system(s)
        one.set_system(s)
        rng = np.random.Generator(np.random.PCG64(11))
        for step in range(4):
            first = 5 * int(rng.integers(0, s.n // 5))
            s.pos[first:first + 5] += rng.uniform(-0.2, 0.2, size=3)
            group.set_positions(first, s.pos[first:first + 5])
            one.set_positions(first, s.pos[first:first + 5])
            a, b = group.energy(), one.energy()
            for k in ("rd", "es_real", "es_recip", "polar"):
                assert abs(getattr(a, k) - getattr(b, k)) <= 1e-11 * abs(getattr(b, k)), (step, k)
        for n in (2000, 5000):
            s = systems.h2_fluid(n)
            group.set_system(s)
            one.set_system(s)
            a, b = group.energy(), one.energy()
            assert abs(a.total - b.total) <= 1e-11 * (abs(b.rd) + abs(b.polar) + abs(b.coulombic))
            err = np.abs(group.atom_arrays()["mu"] - one.atom_arrays()["mu"]).max()
            assert err <= 1e-11 * np.abs(one.atom_arrays()["mu"]).max()
    finally:
        one.close()


def test_group_refuses_what_does_not_shard(group):
    s = systems.h2_fluid(1000)
    s.params.polar_gs = 1
    group.set_system(s)
    with pytest.raises(RuntimeError, match="Gauss-Seidel"):
        group.energy()
    s = systems.h2_fluid(1000)
    s.params.cavity_autoreject_absolute, s.params.cavity_autoreject_scale = 1, 0.5
    group.set_system(s)
    with pytest.raises(RuntimeError, match="cavity_autoreject_absolute"):
        group.energy()
    s = systems.h2_fluid(1000)
    group.set_system(s)
    assert group.energy().total != 0.0  # the group stays usable
