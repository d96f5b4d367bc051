"""World-size-2 gloo tests (CPU) of the multi-GPU host logic in mpmc_b200/sharded.py: the row
bookkeeping, the in-place all-gather of SoA row blocks, and the staged solver protocol
(thole_iterative cut at its exchange points, reference src/polarization/thole_iterative.c:168-259)
driven against a small dense host stand-in for the engine.  No CUDA, no oracle."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mpmc_b200 import api
from mpmc_b200.sharded import Exchange, ShardedEngine, padded_rows, row_block, shard_range

TILE = 128


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def test_row_blocks_partition_the_rows():
    for n in (1, 127, 128, 129, 1000, 46413):
        for world in (1, 2, 4, 8):
            npad = padded_rows(n, world)
            assert npad % (TILE * world) == 0 and npad >= n
            blocks = [row_block(npad, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == npad
            assert all(blocks[r][1] == blocks[r + 1][0] for r in range(world - 1))
            assert len({e - b for b, e in blocks}) == 1  # equal blocks => plain all-gather
    assert [shard_range(10, r, 4) for r in range(4)] == [(0, 2), (2, 5), (5, 7), (7, 10)]


class HostEngine:
    """Dense stand-in with the staged interface of api.Engine: E_ind = -T mu on a random symmetric
    T, mu_new = alpha (E_s + E_ind); owns rows row_block(pp, rank, world)."""

    torch_device = "cpu"

    def __init__(self, n_pol, seed=3, sym=True):
        self.n_pol, self.rank, self.world = n_pol, 0, 1
        self.seed = seed
        self.sym = sym

    def uses_symmetric_contraction(self):
        return self.sym

    def set_shard(self, rank, world):
        self.rank, self.world = rank, world

    def set_system(self, s):
        self.p = s.params
        pp = self.pp = padded_rows(self.n_pol, self.world)
        self.npad = padded_rows(self.n_pol + 7, self.world)
        rng = np.random.default_rng(self.seed)
        n = self.n_pol
        T = rng.normal(size=(3 * n, 3 * n)) * (0.25 / np.sqrt(3 * n))
        self.T = T + T.T
        self.alpha = rng.uniform(0.5, 1.0, size=n)
        self.es_full = rng.normal(size=(3, n))
        self.bufs = {0: torch.zeros(3 * pp, dtype=torch.float64), 2: torch.zeros(32, dtype=torch.float64),
                     3: torch.zeros(3 * self.npad, dtype=torch.float64), 4: torch.zeros(3 * pp, dtype=torch.float64)}
        self.mu_new = np.zeros((3, pp))
        self.mu_old = np.zeros((3, pp))
        self.eind = np.zeros((3, pp))
        self.eic = np.zeros((3, pp))
        self.log = []

    def buffer_tensor(self, which):
        return self.bufs[which]

    def shard_rows(self):
        b, e = row_block(self.pp, self.rank, self.world)
        return b, e, self.n_pol, self.pp

    def shard_field_rows(self):
        b, e = row_block(self.npad, self.rank, self.world)
        return b, e, self.npad

    def energy_enqueue(self):
        self.bufs[2].zero_()

    def _rows(self):
        b, e, n, _ = self.shard_rows()
        return b, min(e, n)

    def _contract(self, b, e):
        mu = self.bufs[0].numpy().reshape(3, self.pp)[:, :self.n_pol]
        x = mu.T.reshape(-1)  # atom-major 3-vectors
        rows = np.arange(3 * b, 3 * e)
        return -(self.T[rows] @ x).reshape(e - b, 3).T

    def polar_stage(self, stage, iteration=0):
        self.log.append(stage)
        mu = self.bufs[0].numpy().reshape(3, self.pp)
        b, e = self._rows()
        if stage == api.STAGE_FIELD:  # "static field" of own atom rows (atoms 0..n_pol-1 are the polarizable ones)
            fb, fe, npad = self.shard_field_rows()
            ef = self.bufs[3].numpy().reshape(3, npad)
            ef[:] = 0.0
            hi = min(fe, self.n_pol)
            if hi > fb:
                ef[:, fb:hi] = self.es_full[:, fb:hi]
        elif stage == api.STAGE_INIT:
            ef = self.bufs[3].numpy().reshape(3, self.npad)
            self.es = ef[:, :self.n_pol].copy()
            mu[:] = 0.0
            mu[:, :self.n_pol] = self.alpha * self.es * self.p.polar_gamma
            self.eic[:] = 0.0
        elif self.sym and stage in (api.STAGE_PASS, api.STAGE_PALMO):
            # this rank's share of the pair sum (its column block of T) for ALL rows -> buffer 4
            n = self.n_pol
            x = mu[:, :n].T.reshape(-1)
            cols = np.arange(3 * b, 3 * e)
            part = -(self.T[:, cols] @ x[cols]).reshape(n, 3).T
            es4 = self.bufs[4].numpy().reshape(3, self.pp)
            es4[:] = 0.0
            es4[:, :n] = part
        elif self.sym and stage == api.STAGE_UPDATE:
            n = self.n_pol
            es4 = self.bufs[4].numpy().reshape(3, self.pp)
            self.mu_old[:, :n] = mu[:, :n]
            self.eind[:, :n] = es4[:, :n]
            self.mu_new[:, :n] = self.alpha * (self.es + self.eind[:, :n])
            d = self.mu_new[:, :n] - self.mu_old[:, :n]
            self.bufs[2][12] = float((d * d).max())
        elif self.sym and stage == api.STAGE_PALMO_FINISH:
            n = self.n_pol
            es4 = self.bufs[4].numpy().reshape(3, self.pp)
            self.eic[:, :n] = -self.eind[:, :n] + es4[:, :n]
        elif self.sym and stage == api.STAGE_COMMIT:
            mu[:, :self.n_pol] = self.mu_new[:, :self.n_pol]
        elif stage in (api.STAGE_UPDATE, api.STAGE_PALMO_FINISH):
            pass
        elif stage == api.STAGE_PASS:
            self.mu_old[:, b:e] = mu[:, b:e]
            self.eind[:, b:e] = self._contract(b, e)
            self.mu_new[:, b:e] = self.alpha[b:e] * (self.es[:, b:e] + self.eind[:, b:e])
            d = self.mu_new[:, b:e] - self.mu_old[:, b:e]
            self.bufs[2][12] = float((d * d).max()) if e > b else 0.0
        elif stage == api.STAGE_PALMO:
            self.eic[:, b:e] = -self.eind[:, b:e] + self._contract(b, e)
        elif stage == api.STAGE_COMMIT:
            mu[:, b:e] = self.mu_new[:, b:e]
        elif stage == api.STAGE_FAIL:
            mu[:, :self.n_pol] = self.alpha * self.es
        elif stage == api.STAGE_ENERGY:
            self.iterations = iteration
            pot = (mu[:, b:e] * self.es[:, b:e]).sum()
            if self.p.polar_palmo:
                pot += (mu[:, b:e] * self.eic[:, b:e]).sum()
            self.upol = -0.5 * pot

    def energy_finish(self):
        e = api.Energies()
        e.polar = self.upol
        e.total = self.upol
        return e

    def counters(self):
        return api.Counters()

    def atom_arrays(self):
        mu = self.bufs[0].numpy().reshape(3, self.pp)[:, :self.n_pol].T.copy()
        own = np.zeros((3, self.pp))
        b, e = self._rows()
        own[:, b:e] = 1.0  # like the engine: summed arrays are reported for own rows only
        return {"mu": mu, "ef_static": self.es.T.copy(), "ef_induced": (self.eind * own)[:, :self.n_pol].T.copy(),
                "ef_induced_change": (self.eic * own)[:, :self.n_pol].T.copy(), "dipole_rrms": np.zeros(self.n_pol)}


def _system(**kw):
    p = api.default_params(np.eye(3) * 30.0)
    p.polarization, p.polar_iterative = 1, 1
    for k, v in kw.items():
        setattr(p, k, v)
    z = np.zeros(1)
    return api.System(p, np.zeros((1, 3)), z, z, z, z, z, np.zeros(1, dtype=np.int32), np.zeros(1, dtype=np.int32))


def _solve(rank, world, n_pol, kw, sym=True):
    eng = HostEngine(n_pol, sym=sym)
    sh = ShardedEngine(eng, rank, world)
    sh.set_system(_system(**kw))
    e = sh._energy_sharded()
    return e, sh.atom_arrays(), eng.log


def _worker(rank, world, port, n_pol, kw, ret, sym=True):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # 1. in-place all-gather of SoA row blocks
        npad = padded_rows(1000, world)
        b, e = row_block(npad, rank, world)
        buf = torch.full((3 * npad,), -1.0, dtype=torch.float64)
        full = torch.arange(3 * npad, dtype=torch.float64)
        for c in range(3):
            buf[c * npad + b:c * npad + e] = full[c * npad + b:c * npad + e]
        Exchange(rank, world).all_gather_rows(buf, 3, npad)
        assert torch.equal(buf, full)
        # 2. the staged solver
        e, arr, log = _solve(rank, world, n_pol, kw, sym)
        ret[rank] = (e.polar, e.polar_iterations, e.iter_success, arr["mu"], arr["ef_induced"], arr["ef_induced_change"], log)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("kw", [dict(polar_max_iter=6), dict(polar_max_iter=4, polar_palmo=1),
                                dict(polar_max_iter=0, polar_precision=1e-9),
                                dict(polar_max_iter=0, polar_precision=1e-9, polar_palmo=1)])
@pytest.mark.parametrize("sym", [True, False])
def test_two_rank_solver_equals_one_rank(kw, sym):
    """sym=True: symmetric contraction, all-reduce of the induced field; sym=False: row form,
    all-gather of the dipole rows (MPMC_B200_K4=rows)."""
    n_pol, world = 300, 2
    ref_e, ref_arr, ref_log = _solve(0, 1, n_pol, kw, sym)
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), n_pol, kw, ret, sym), nprocs=world, join=True)
    assert ref_e.polar_iterations > 1
    for r in range(world):
        upol, it, fail, mu, eind, eic, log = ret[r]
        assert it == ref_e.polar_iterations and fail == ref_e.iter_success
        assert log == ref_log  # same stage sequence on every rank
        assert abs(upol - ref_e.polar) <= 1e-12 * abs(ref_e.polar)  # summed over ranks
        np.testing.assert_allclose(mu, ref_arr["mu"], rtol=0, atol=1e-13)
        np.testing.assert_allclose(eind, ref_arr["ef_induced"], rtol=0, atol=1e-13)
        np.testing.assert_allclose(eic, ref_arr["ef_induced_change"], rtol=0, atol=1e-13)
    if kw.get("polar_palmo"):
        assert ref_log.count(api.STAGE_PALMO) == 1 and ref_log[-4 if sym else -3] == api.STAGE_PALMO
