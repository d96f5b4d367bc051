"""Row-sharded evaluation of energy() across the GPUs of one NVSwitch box: one process per GPU,
``torch.distributed`` (NCCL over NVLink) for the exchanges, the CUDA engine for all arithmetic.

What is split (BASELINE north_star):
  * pair tiles of the LJ + Ewald-real kernel and the atom chunks of the structure factor
    -> partial scalars, summed over ranks at the end;
  * static-field rows (thole_field, reference src/polarization/thole_field.c:14-36), by spatial cluster
    -> one all-reduce(sum) of the field (each row is written by exactly one rank, zero elsewhere);
  * the tile pairs of every solver pass (contract_dipoles, src/polarization/thole_iterative.c:27-59):
    the dipole tensor is symmetric, so each rank evaluates T_ij once for its share of the unordered
    tile pairs and applies it in both directions -> its contribution to the induced field of ALL rows
    -> an all-reduce(sum) of the induced field (3 n_pol doubles) after EVERY pass: the one real
       exchange of the path.  Every rank then updates every dipole (a few microseconds), so no
       dipole all-gather and no residual all-reduce are needed.
    (With the row-by-row contraction, MPMC_B200_K4=rows, rows are owned instead: an all-gather of
    the new dipole rows after every pass plus an all-reduce(max) of the residual.)
Gauss-Seidel modes are sequential in atom order and stay on one GPU.

Default since round 2: once the ranks have exchanged the CUDA-IPC handles of their exchange buffers (``_connect_peers``,
the only use of ``torch.distributed`` on this path) ``energy()`` is ONE call into the library -- the staged loop below
lives in ``engine.cu:sharded_eval`` and every exchange is a kernel of the engine reading peer memory.  The Python staging
with NCCL collectives remains as the fallback / A-B switch (``MPMC_B200_PEER=0`` or ``MPMC_B200_STAGED=1``) and is what
the gloo tests exercise on the CPU.

The collectives are issued by ``Exchange`` on the same CUDA stream as the engine's kernels, so no
host synchronisation separates a pass from its exchange.  ``Exchange`` only needs ``all_gather``
/ ``all_reduce`` on flat buffers, so the row bookkeeping is testable on CPU with gloo
(tests/test_sharded_cpu.py).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import numpy as np

from . import api
from .api import Counters, Energies

MAX_ITERATION_COUNT = 128  # reference src/include/defines.h:MAX_ITERATION_COUNT
DEBYE2SKA = api.DEBYE2SKA
TILE = 128

_ENERGY_SUM_FIELDS = ("total", "rd", "coulombic", "polar", "rd_pair", "rd_lrc", "es_real", "es_recip", "es_self",
                      "dipole_rrms")
_COUNTER_SUM_FIELDS = ("pairs_tested", "pairs_lj", "pairs_es", "pairs_field")


def shard_range(units: int, rank: int, world: int):
    """[begin, end) of `units` work units owned by `rank` -- same formula as the engine's shard_range()."""
    return units * rank // world, units * (rank + 1) // world


def padded_rows(n: int, world: int, tile: int = TILE) -> int:
    m = tile * max(world, 1)
    return max(((n + m - 1) // m) * m, tile)


def row_block(n_pad: int, rank: int, world: int, tile: int = TILE):
    """Tile-aligned row block of a padded SoA array; equal on every rank because n_pad is a
    multiple of tile*world (that is what makes the all-gather a plain NCCL all-gather)."""
    tb, te = shard_range(n_pad // tile, rank, world)
    return tb * tile, te * tile


class Exchange:
    """The collectives of the sharded path on flat SoA buffers ``buf[c*n_pad + row]``."""

    def __init__(self, rank: int, world: int, group=None):
        self.rank, self.world, self.group = rank, world, group

    def all_gather_rows(self, buf, ncomp: int, n_pad: int):
        """buf: 1-D torch tensor [ncomp*n_pad] whose rows row_block(n_pad, rank) are valid on
        `rank`; afterwards every rank holds every row.  In place (input is a view of output)."""
        if self.world == 1:
            return
        import torch.distributed as dist

        b, e = row_block(n_pad, self.rank, self.world)
        for c in range(ncomp):
            comp = buf[c * n_pad:(c + 1) * n_pad]
            dist.all_gather_into_tensor(comp, comp[b:e], group=self.group)

    def all_reduce(self, t, op: str = "sum"):
        if self.world == 1:
            return
        import torch.distributed as dist

        dist.all_reduce(t, op=dist.ReduceOp.SUM if op == "sum" else dist.ReduceOp.MAX, group=self.group)


class _DevView:
    """Minimal __cuda_array_interface__ holder so a raw engine pointer becomes a torch tensor
    (no copy; the engine owns the memory)."""

    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (ptr, False), "version": 3}


@dataclass
class OwnCounters:
    pairs_dipole_own: float = 0.0


class ShardedEngine:
    """energy() of one system over `world` GPUs.  world == 1 is a plain pass-through to Engine."""

    def __init__(self, engine: api.Engine, rank: int = 0, world: int = 1, stream=None, group=None):
        self.eng, self.rank, self.world = engine, rank, world
        self.ex = Exchange(rank, world, group)
        self.stream = stream
        self.device = getattr(engine, "torch_device", "cuda")
        self.system = None
        self._launches = 0
        self._last_cnt: Optional[Counters] = None
        self._views = {}

    # -- system ------------------------------------------------------------
    def set_system(self, s: api.System):
        self.system = s
        if self.world > 1:
            p = s.params
            if p.polarization and (p.polar_gs or p.polar_gs_ranked):
                raise api.EngineError("Gauss-Seidel polarization is sequential in atom order: single-GPU only")
            if p.cavity_autoreject_absolute:
                raise api.EngineError("cavity_autoreject_absolute is evaluated on one GPU only")
            self.eng.set_shard(self.rank, self.world)
        self.eng.set_system(s)
        self._views = {}
        self._connect_peers()

    def _connect_peers(self):
        """Exchange the CUDA IPC handles of the ranks' exchange buffers (mpmc_b200_peer_export /
        peer_connect): afterwards the per-iteration all-reduce of the induced field is done by the
        engine's own update kernel over NVLink peer memory.  MPMC_B200_PEER=0 keeps the NCCL all-reduce."""
        import os

        eng = self.eng
        if self.world == 1 or not hasattr(eng, "peer_export") or os.environ.get("MPMC_B200_PEER", "1") == "0":
            return
        if not eng.uses_symmetric_contraction() or self.world > 8:
            return
        import torch.distributed as dist

        handles = [None] * self.world
        dist.all_gather_object(handles, eng.peer_export(), group=self.ex.group)
        eng.peer_connect(handles)
        dist.barrier(group=self.ex.group)

    def set_positions(self, first: int, pos: np.ndarray):
        self.eng.set_positions(first, pos)

    def _view(self, which: int):
        import torch

        if which not in self._views:
            if hasattr(self.eng, "buffer_tensor"):  # host stand-in used by the gloo tests
                self._views[which] = self.eng.buffer_tensor(which)
            else:
                ptr, n = self.eng.device_buffer(which)
                self._views[which] = torch.as_tensor(_DevView(ptr, n), device="cuda")
        return self._views[which]

    # -- evaluation ----------------------------------------------------------
    def energy(self) -> Energies:
        if self.world == 1:
            e = self.eng.energy()
            self._last_cnt = self.eng.counters()
            self._launches = self._last_cnt.kernel_launches
            return e
        return self._energy_sharded()

    def _energy_sharded(self) -> Energies:
        import os

        import torch

        eng, p = self.eng, self.system.params
        if bool(getattr(eng, "peer_active", lambda: False)()) and os.environ.get("MPMC_B200_STAGED", "0") != "1":
            # The whole staged evaluation inside the library (engine.cu: sharded_eval): static-field rows, the induced
            # field of every pass and the final scalars are exchanged by the engine's own kernels over NVLink peer
            # memory; energies and counters come back as totals, identical on every rank.  No NCCL call at all.
            out = eng.energy()
            self._last_cnt = eng.counters()
            self._launches = self._last_cnt.kernel_launches
            return out
        eng.energy_enqueue()
        iterations, failed = 0, False
        polar = bool(p.polarization and not p.rd_only)
        if polar:
            _, _, n_pol, pp = eng.shard_rows()
            _, _, np_atoms = eng.shard_field_rows()
            eng.polar_stage(api.STAGE_FIELD)
            if getattr(eng, "uses_cluster_field", lambda: False)():
                self.ex.all_reduce(self._view(3), "sum")  # rows of other ranks' clusters are zero here
            else:
                self.ex.all_gather_rows(self._view(3), 3, np_atoms)
            eng.polar_stage(api.STAGE_INIT)
            mu = self._view(0)
            sym = bool(eng.uses_symmetric_contraction())
            peer = sym and bool(getattr(eng, "peer_active", lambda: False)())  # exchange inside the update kernel
            if not p.polar_zodid and n_pol > 0:
                allowed = p.polar_precision * p.polar_precision * DEBYE2SKA * DEBYE2SKA  # thole_iterative.c:106
                keep = True
                while keep:
                    iterations += 1
                    if iterations >= MAX_ITERATION_COUNT and p.polar_precision != 0.0:
                        eng.polar_stage(api.STAGE_FAIL)
                        failed = True
                        break
                    eng.polar_stage(api.STAGE_PASS, iterations)
                    if sym:
                        if not peer:
                            self.ex.all_reduce(self._view(4), "sum")
                        eng.polar_stage(api.STAGE_UPDATE, iterations)
                    if p.polar_precision == 0.0:
                        keep = iterations != p.polar_max_iter
                    else:
                        err = self._view(2)[12:13]
                        if not sym:
                            self.ex.all_reduce(err, "max")
                        keep = float(err.item()) > allowed
                    if p.polar_palmo and not keep:
                        eng.polar_stage(api.STAGE_PALMO)
                        if sym:
                            if not peer:
                                self.ex.all_reduce(self._view(4), "sum")
                            eng.polar_stage(api.STAGE_PALMO_FINISH)
                    eng.polar_stage(api.STAGE_COMMIT, iterations)
                    if not sym:
                        self.ex.all_gather_rows(mu, 3, pp)
            eng.polar_stage(api.STAGE_ENERGY, iterations)
        part = eng.energy_finish()
        cnt = eng.counters()
        self._launches = cnt.kernel_launches
        vals = [getattr(part, k) for k in _ENERGY_SUM_FIELDS] + [getattr(cnt, k) for k in _COUNTER_SUM_FIELDS]
        t = torch.tensor(vals, dtype=torch.float64, device=self.device)
        self.ex.all_reduce(t, "sum")
        vals = t.tolist()
        out = Energies()
        for k, v in zip(_ENERGY_SUM_FIELDS, vals):
            setattr(out, k, v)
        for k, v in zip(_COUNTER_SUM_FIELDS, vals[len(_ENERGY_SUM_FIELDS):]):
            setattr(cnt, k, v)
        out.polar_iterations = iterations
        out.iter_success = 1 if failed else 0
        out.min_intermolecular_r = part.min_intermolecular_r
        self._last_cnt = cnt
        return out

    def atom_arrays(self, out: Optional[dict] = None) -> dict:
        arr = self.eng.atom_arrays(out) if out is not None else self.eng.atom_arrays()
        if self.world > 1:
            import torch

            complete = getattr(self.eng, "atom_arrays_complete", lambda: False)()
            if complete:
                # symmetric contraction: every rank holds every row; only the palmo rows of the
                # non-polarizable atoms (write-back only) are computed per rank
                if self.system.params.polar_palmo:
                    eic = arr["ef_induced_change"]
                    if self.rank != 0:
                        eic[np.asarray(self.system.polarizability) != 0.0] = 0.0
                    t = torch.from_numpy(eic).to(self.device)
                    self.ex.all_reduce(t, "sum")
                    arr["ef_induced_change"] = t.cpu().numpy()
                return arr
            # row form: mu and ef_static are complete on every rank; the rest is reported for the
            # owning rank's rows and zero elsewhere -> sum over ranks
            for k in ("ef_induced", "ef_induced_change", "dipole_rrms"):
                t = torch.from_numpy(arr[k]).to(self.device)
                self.ex.all_reduce(t, "sum")
                arr[k] = t.cpu().numpy()
        return arr

    # -- bookkeeping -----------------------------------------------------------
    def kernel_launches(self) -> int:
        return int(self._launches)

    def counters_total(self) -> Counters:
        return self._last_cnt

    def counters_own(self) -> OwnCounters:
        c = self._last_cnt
        b, e, n_pol, _ = self.eng.shard_rows() if c.n_polarizable > 0 else (0, 0, 0, 0)
        if self.world > 1 and self.eng.uses_symmetric_contraction():
            # symmetric K4: the (I tile, offset slice) work items are dealt out evenly, 1/world of the pairs each
            return OwnCounters(pairs_dipole_own=float(n_pol) * max(n_pol - 1, 0) / self.world)
        rows = max(0, min(e, n_pol) - b)
        return OwnCounters(pairs_dipole_own=float(rows) * max(n_pol - 1, 0))
