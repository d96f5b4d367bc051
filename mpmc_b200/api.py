"""Host-side mirror of the reference's energy interface on top of the C ABI.

``Engine.energy()`` / ``Engine.polar()`` have the meaning of the reference's
``double energy(system_t*)`` (src/energy/energy.c:67-226) and
``double polar(system_t*)`` (src/energy/polar.c:31-135); ``System`` is the flat
(array) form of the ``system_t -> molecule_t -> atom_t`` lists in the order
``rebuild_arrays()`` produces (src/energy/pairs.c:12-38); ``Params`` carries the
``system_t``/``pbc_t`` scalars the hot path reads (src/include/structs.h).

Everything here goes through ``libmpmc_b200.so`` (include/mpmc_b200.h).  There
is no CPU fallback: if the CUDA library is missing or no device is present the
constructors raise.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MPMC_B200_LIB") or os.path.join(_HERE, "lib", "libmpmc_b200.so")  # env: profiling builds

DAMP_OFF, DAMP_LINEAR, DAMP_EXPONENTIAL = 0, 1, 2
MIX_LB, MIX_WH, MIX_HALGREN, MIX_C6 = 0, 1, 2, 3

# enum of mpmc_b200_polar_stage (include/mpmc_b200.h)
(STAGE_FIELD, STAGE_INIT, STAGE_PASS, STAGE_PALMO, STAGE_COMMIT, STAGE_ENERGY, STAGE_FAIL, STAGE_UPDATE,
 STAGE_PALMO_FINISH) = range(9)

E2REDUCED = 408.7816  # reference src/include/defines.h:45
DEBYE2SKA = 85.10597636  # defines.h:51


class Params(C.Structure):
    """struct mpmc_b200_params (include/mpmc_b200.h)."""

    _fields_ = [
        ("basis", (C.c_double * 3) * 3),
        ("reciprocal_basis", (C.c_double * 3) * 3),
        ("cutoff", C.c_double),
        ("volume", C.c_double),
        ("ewald_alpha", C.c_double),
        ("ewald_kmax", C.c_int),
        ("feynman_hibbs", C.c_int),
        ("feynman_hibbs_order", C.c_int),
        ("temperature", C.c_double),
        ("rd_lrc", C.c_int),
        ("mixing", C.c_int),
        ("rd_only", C.c_int),
        ("cavity_autoreject_absolute", C.c_int),
        ("cavity_autoreject_scale", C.c_double),
        ("polarization", C.c_int),
        ("polar_iterative", C.c_int),
        ("damp_type", C.c_int),
        ("polar_damp", C.c_double),
        ("polar_max_iter", C.c_int),
        ("polar_precision", C.c_double),
        ("polar_gamma", C.c_double),
        ("polar_gs", C.c_int),
        ("polar_gs_ranked", C.c_int),
        ("polar_sor", C.c_int),
        ("polar_esor", C.c_int),
        ("polar_palmo", C.c_int),
        ("polar_rrms", C.c_int),
        ("polar_zodid", C.c_int),
        ("polar_ewald", C.c_int),
        ("polar_ewald_full", C.c_int),
        ("polar_wolf", C.c_int),
        ("polar_wolf_full", C.c_int),
        ("polar_wolf_alpha_lookup", C.c_int),
        ("polar_wolf_alpha", C.c_double),
        ("polar_wolf_alpha_lookup_cutoff", C.c_double),
        ("polar_ewald_alpha", C.c_double),
        ("es_wolf", C.c_int),
        ("rd_crystal", C.c_int),
        ("rd_crystal_order", C.c_int),
        ("reserved0", C.c_int),
    ]

    def to_dict(self) -> dict:
        d = {}
        for name, _ in self._fields_:
            v = getattr(self, name)
            if name in ("basis", "reciprocal_basis"):
                v = [[v[i][j] for j in range(3)] for i in range(3)]
            d[name] = v
        return d

    @classmethod
    def from_dict(cls, d: dict) -> "Params":
        p = cls()
        for name, _ in cls._fields_:
            if name not in d:
                continue
            v = d[name]
            if name in ("basis", "reciprocal_basis"):
                m = np.asarray(v, dtype=np.float64).reshape(3, 3)
                for i in range(3):
                    for j in range(3):
                        getattr(p, name)[i][j] = float(m[i, j])
            else:
                setattr(p, name, v)
        return p


class Energies(C.Structure):
    """struct mpmc_b200_energies."""

    _fields_ = [
        ("total", C.c_double),
        ("rd", C.c_double),
        ("coulombic", C.c_double),
        ("polar", C.c_double),
        ("rd_pair", C.c_double),
        ("rd_lrc", C.c_double),
        ("es_real", C.c_double),
        ("es_recip", C.c_double),
        ("es_self", C.c_double),
        ("dipole_rrms", C.c_double),
        ("min_intermolecular_r", C.c_double),
        ("polar_iterations", C.c_int),
        ("iter_success", C.c_int),
        ("autorejected", C.c_int),
        ("reserved", C.c_int),
    ]

    def to_dict(self) -> dict:
        return {n: getattr(self, n) for n, _ in self._fields_ if n != "reserved"}


class Counters(C.Structure):
    """struct mpmc_b200_counters."""

    _fields_ = [
        ("pairs_tested", C.c_double),
        ("pairs_lj", C.c_double),
        ("pairs_es", C.c_double),
        ("pairs_field", C.c_double),
        ("pairs_dipole", C.c_double),
        ("contractions", C.c_double),
        ("kvectors", C.c_double),
        ("recip_atoms", C.c_double),
        ("n_atoms", C.c_int),
        ("n_polarizable", C.c_int),
        ("kernel_launches", C.c_int),
        ("incremental", C.c_int),
        ("pairs_induced_list", C.c_double),
    ]

    def to_dict(self) -> dict:
        return {n: getattr(self, n) for n, _ in self._fields_ if n != "reserved"}


MC_MAX_ATOMS, MC_MAX_STEPS = 16, 64  # MPMC_B200_MC_MAX_ATOMS / MPMC_B200_MC_MAX_STEPS


class McMove(C.Structure):
    """struct mpmc_b200_mc_move (resident MC steps, mpmc_b200_mc_steps)."""

    _fields_ = [("first", C.c_int), ("count", C.c_int), ("explicit_positions", C.c_int), ("reserved", C.c_int),
                ("u_accept", C.c_double), ("trans", C.c_double * 3), ("quat", C.c_double * 4), ("mass", C.c_double * MC_MAX_ATOMS),
                ("new_pos", C.c_double * (3 * MC_MAX_ATOMS)), ("old_pos", C.c_double * (3 * MC_MAX_ATOMS))]


class McResult(C.Structure):
    """struct mpmc_b200_mc_result."""

    _fields_ = [("e", Energies), ("accepted", C.c_int), ("reserved", C.c_int), ("new_pos", C.c_double * (3 * MC_MAX_ATOMS)),
                ("old_pos", C.c_double * (3 * MC_MAX_ATOMS))]


def default_params(basis, cutoff: Optional[float] = None) -> Params:
    """Defaults of the reference's ``setdefaults()`` (src/io/input.c:1598-1667) and
    ``pbc()`` (src/energy/pbc.c:65-83): kmax 7, gamma 1, max_iter 10, rd_lrc on,
    cutoff = half the shortest lattice vector, ewald_alpha = 3.5/cutoff."""
    p = Params()
    b = np.asarray(basis, dtype=np.float64).reshape(3, 3)
    vol = float(np.linalg.det(b))
    # pbc_reciprocal(): reciprocal_basis = inverse(basis)^T in the [q][p] convention used by minimum_image
    rb = _reciprocal(b)
    for i in range(3):
        for j in range(3):
            p.basis[i][j] = float(b[i, j])
            p.reciprocal_basis[i][j] = float(rb[i, j])
    p.volume = _volume(b)
    assert abs(p.volume - vol) <= 1e-9 * max(1.0, abs(vol))
    p.cutoff = float(cutoff) if cutoff else pbc_cutoff(b)
    p.ewald_alpha = 3.5 / p.cutoff
    p.polar_ewald_alpha = 3.5 / p.cutoff
    p.ewald_kmax = 7
    p.polar_gamma = 1.0
    p.polar_max_iter = 10
    p.rd_lrc = 1
    p.temperature = 0.0
    p.polar_wolf_alpha_lookup_cutoff = 30.0
    return p


def _volume(b: np.ndarray) -> float:
    """pbc_volume(): src/energy/pbc.c:36-44 (same expression order)."""
    v = b[0, 0] * (b[1, 1] * b[2, 2] - b[1, 2] * b[2, 1])
    v += b[0, 1] * (b[1, 2] * b[2, 0] - b[1, 0] * b[2, 2])
    v += b[0, 2] * (b[1, 0] * b[2, 1] - b[1, 1] * b[2, 0])
    return float(v)


def _reciprocal(b: np.ndarray) -> np.ndarray:
    """pbc_reciprocal(): src/energy/pbc.c:47-63 (same expression order)."""
    iv = 1.0 / _volume(b)
    r = np.zeros((3, 3))
    r[0, 0] = iv * (b[1, 1] * b[2, 2] - b[1, 2] * b[2, 1])
    r[0, 1] = iv * (b[0, 2] * b[2, 1] - b[0, 1] * b[2, 2])
    r[0, 2] = iv * (b[0, 1] * b[1, 2] - b[0, 2] * b[1, 1])
    r[1, 0] = iv * (b[1, 2] * b[2, 0] - b[1, 0] * b[2, 2])
    r[1, 1] = iv * (b[0, 0] * b[2, 2] - b[0, 2] * b[2, 0])
    r[1, 2] = iv * (b[0, 2] * b[1, 0] - b[0, 0] * b[1, 2])
    r[2, 0] = iv * (b[1, 0] * b[2, 1] - b[1, 1] * b[2, 0])
    r[2, 1] = iv * (b[0, 1] * b[2, 0] - b[0, 0] * b[2, 1])
    r[2, 2] = iv * (b[0, 0] * b[1, 1] - b[0, 1] * b[1, 0])
    return r


def pbc_cutoff(b: np.ndarray) -> float:
    """pbc_cutoff(): src/energy/pbc.c:13-33 -- half the shortest lattice vector."""
    rng = np.arange(-5, 6)
    i, j, k = np.meshgrid(rng, rng, rng, indexing="ij")
    coef = np.stack([i.ravel(), j.ravel(), k.ravel()], axis=1)
    coef = coef[np.any(coef != 0, axis=1)]
    vec = coef @ b
    return float(0.5 * np.sqrt((vec * vec).sum(axis=1)).min())


@dataclass
class System:
    """Flat form of the reference's molecule/atom lists (list order)."""

    params: Params
    pos: np.ndarray  # [n,3]
    charge: np.ndarray
    polarizability: np.ndarray
    epsilon: np.ndarray
    sigma: np.ndarray
    mol_mass: np.ndarray
    mol_id: np.ndarray  # int32
    frozen: np.ndarray  # int32
    name: str = ""
    meta: dict = field(default_factory=dict)

    def __post_init__(self):
        f8 = lambda a: np.ascontiguousarray(a, dtype=np.float64)
        self.pos = f8(self.pos).reshape(-1, 3)
        n = self.pos.shape[0]
        for k in ("charge", "polarizability", "epsilon", "sigma", "mol_mass"):
            a = f8(getattr(self, k)).reshape(-1)
            assert a.shape[0] == n, k
            setattr(self, k, a)
        self.mol_id = np.ascontiguousarray(self.mol_id, dtype=np.int32).reshape(-1)
        self.frozen = np.ascontiguousarray(self.frozen, dtype=np.int32).reshape(-1)
        assert self.mol_id.shape[0] == n and self.frozen.shape[0] == n

    @property
    def n(self) -> int:
        return int(self.pos.shape[0])

    def copy(self) -> "System":
        p = Params.from_dict(self.params.to_dict())
        return System(p, self.pos.copy(), self.charge.copy(), self.polarizability.copy(), self.epsilon.copy(),
                      self.sigma.copy(), self.mol_mass.copy(), self.mol_id.copy(), self.frozen.copy(), self.name,
                      dict(self.meta))

    def save_arrays(self) -> dict:
        import json

        return dict(params_json=np.array(json.dumps(self.params.to_dict())), pos=self.pos, charge=self.charge,
                    polarizability=self.polarizability, epsilon=self.epsilon, sigma=self.sigma, mol_mass=self.mol_mass,
                    mol_id=self.mol_id, frozen=self.frozen)

    @classmethod
    def from_arrays(cls, z, name="") -> "System":
        import json

        p = Params.from_dict(json.loads(str(z["params_json"])))
        return cls(p, z["pos"], z["charge"], z["polarizability"], z["epsilon"], z["sigma"], z["mol_mass"], z["mol_id"],
                   z["frozen"], name)


def _dp(a: Optional[np.ndarray]):
    if a is None:
        return None
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_int))


_lib = None


def load_library(path: str = LIB_PATH):
    """dlopen libmpmc_b200.so and declare every prototype of include/mpmc_b200.h."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(path):
        raise RuntimeError(
            f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback)")
    lib = C.CDLL(path)
    dp, ip, vp = C.POINTER(C.c_double), C.POINTER(C.c_int), C.c_void_p
    lib.mpmc_b200_abi_version.restype = C.c_int
    lib.mpmc_b200_last_error.restype = C.c_char_p
    lib.mpmc_b200_create.argtypes = [C.POINTER(vp), C.c_int]
    lib.mpmc_b200_destroy.argtypes = [vp]
    lib.mpmc_b200_destroy.restype = None
    lib.mpmc_b200_create_group.argtypes = [C.POINTER(vp), C.c_int, ip]
    lib.mpmc_b200_group_size.argtypes = [vp]
    lib.mpmc_b200_set_stream.argtypes = [vp, vp]
    lib.mpmc_b200_set_params.argtypes = [vp, C.POINTER(Params)]
    lib.mpmc_b200_set_atoms.argtypes = [vp, C.c_int, dp, dp, dp, dp, dp, dp, ip, ip]
    lib.mpmc_b200_set_positions.argtypes = [vp, C.c_int, C.c_int, dp]
    lib.mpmc_b200_set_shard.argtypes = [vp, C.c_int, C.c_int]
    lib.mpmc_b200_set_incremental.argtypes = [vp, C.c_int]
    lib.mpmc_b200_set_warm_start.argtypes = [vp, C.c_int]
    lib.mpmc_b200_energy.argtypes = [vp, C.POINTER(Energies)]
    lib.mpmc_b200_polar.argtypes = [vp, C.POINTER(Energies)]
    lib.mpmc_b200_energy_enqueue.argtypes = [vp]
    lib.mpmc_b200_energy_finish.argtypes = [vp, C.POINTER(Energies)]
    lib.mpmc_b200_polar_stage.argtypes = [vp, C.c_int, C.c_int]
    lib.mpmc_b200_device_buffer.argtypes = [vp, C.c_int, C.POINTER(vp), C.POINTER(C.c_longlong)]
    lib.mpmc_b200_uses_symmetric_contraction.argtypes = [vp]
    lib.mpmc_b200_uses_cluster_field.argtypes = [vp]
    lib.mpmc_b200_shard_rows.argtypes = [vp, ip, ip, ip, ip]
    lib.mpmc_b200_shard_field_rows.argtypes = [vp, ip, ip, ip]
    lib.mpmc_b200_get_atom_arrays.argtypes = [vp, dp, dp, dp, dp, dp, dp]
    lib.mpmc_b200_atom_arrays_complete.argtypes = [vp]
    lib.mpmc_b200_atom_arrays_view.argtypes = [vp] + [C.POINTER(dp)] * 6
    lib.mpmc_b200_peer_export.argtypes = [vp, C.c_char_p]
    lib.mpmc_b200_peer_connect.argtypes = [vp, C.c_int, C.c_char_p]
    lib.mpmc_b200_peer_active.argtypes = [vp]
    lib.mpmc_b200_get_counters.argtypes = [vp, C.POINTER(Counters)]
    lib.mpmc_b200_phase_ms.argtypes = [vp, C.c_char_p, dp]
    lib.mpmc_b200_measure_fp64_peak.argtypes = [vp, dp]
    for name in ABI_SYMBOLS:
        f = getattr(lib, name)
        if name not in ("mpmc_b200_last_error", "mpmc_b200_destroy"):
            f.restype = C.c_int
    if lib.mpmc_b200_abi_version() != 3:
        raise RuntimeError("libmpmc_b200.so ABI version mismatch")
    _lib = lib
    return lib


ABI_SYMBOLS = [
    "mpmc_b200_abi_version", "mpmc_b200_last_error", "mpmc_b200_create", "mpmc_b200_destroy", "mpmc_b200_set_stream",
    "mpmc_b200_set_params", "mpmc_b200_set_atoms", "mpmc_b200_set_positions", "mpmc_b200_set_shard",
    "mpmc_b200_energy", "mpmc_b200_polar", "mpmc_b200_energy_enqueue", "mpmc_b200_energy_finish",
    "mpmc_b200_polar_stage", "mpmc_b200_uses_symmetric_contraction", "mpmc_b200_uses_cluster_field",
    "mpmc_b200_device_buffer", "mpmc_b200_shard_rows", "mpmc_b200_shard_field_rows",
    "mpmc_b200_get_atom_arrays", "mpmc_b200_atom_arrays_complete",
    "mpmc_b200_peer_export", "mpmc_b200_peer_connect", "mpmc_b200_peer_active",
    "mpmc_b200_get_counters", "mpmc_b200_phase_ms", "mpmc_b200_measure_fp64_peak", "mpmc_b200_set_incremental",
    "mpmc_b200_set_warm_start", "mpmc_b200_create_group", "mpmc_b200_group_size", "mpmc_b200_atom_arrays_view",
    "mpmc_b200_recip_energy_static", "mpmc_b200_mc_steps", "mpmc_b200_mc_step_arrays_view",
]


class EngineError(RuntimeError):
    pass


class Engine:
    """One CUDA context of the energy engine (one per GPU / per MC chain)."""

    def __init__(self, device: int = 0, stream: Optional[int] = None, devices=None):
        """devices: a list of CUDA ordinals makes this the leader of a group -- one process driving several GPUs of one
        NVLink box (mpmc_b200_create_group); every call below then addresses the whole group."""
        self.lib = load_library()
        self._ctx = C.c_void_p()
        if devices is not None:
            d = (C.c_int * len(devices))(*[int(x) for x in devices])
            self._check(self.lib.mpmc_b200_create_group(C.byref(self._ctx), len(devices), d))
        else:
            self._check(self.lib.mpmc_b200_create(C.byref(self._ctx), int(device)))
        self.n = 0
        if stream is not None:
            self._check(self.lib.mpmc_b200_set_stream(self._ctx, C.c_void_p(stream)))

    def _check(self, rc: int):
        if rc != 0:
            msg = self.lib.mpmc_b200_last_error()
            raise EngineError(f"libmpmc_b200 error {rc}: {msg.decode() if msg else '?'}")

    def close(self):
        if getattr(self, "_ctx", None) is not None and self._ctx.value:
            self.lib.mpmc_b200_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- system definition -------------------------------------------------
    def set_params(self, p: Params):
        self._check(self.lib.mpmc_b200_set_params(self._ctx, C.byref(p)))

    def set_system(self, s: System):
        self.set_params(s.params)
        self.n = s.n
        self._check(self.lib.mpmc_b200_set_atoms(self._ctx, s.n, _dp(s.pos), _dp(s.charge), _dp(s.polarizability),
                                                 _dp(s.epsilon), _dp(s.sigma), _dp(s.mol_mass), _ip(s.mol_id),
                                                 _ip(s.frozen)))

    def set_positions(self, first: int, pos: np.ndarray):
        pos = np.ascontiguousarray(pos, dtype=np.float64).reshape(-1, 3)
        self._check(self.lib.mpmc_b200_set_positions(self._ctx, int(first), int(pos.shape[0]), _dp(pos)))

    def set_incremental(self, on: bool):
        """Opt-in incremental energy() after small moves (include/mpmc_b200.h: mpmc_b200_set_incremental)."""
        self._check(self.lib.mpmc_b200_set_incremental(self._ctx, 1 if on else 0))

    def set_warm_start(self, on: bool):
        """Opt-in warm start of the tolerance-mode dipole solve (include/mpmc_b200.h: mpmc_b200_set_warm_start)."""
        self._check(self.lib.mpmc_b200_set_warm_start(self._ctx, 1 if on else 0))

    def last_eval_incremental(self) -> bool:
        return bool(self.counters().incremental)

    def group_size(self) -> int:
        return int(self.lib.mpmc_b200_group_size(self._ctx))

    def set_shard(self, rank: int, world: int):
        self._check(self.lib.mpmc_b200_set_shard(self._ctx, rank, world))

    # -- evaluation --------------------------------------------------------
    def energy(self) -> Energies:
        e = Energies()
        self._check(self.lib.mpmc_b200_energy(self._ctx, C.byref(e)))
        return e

    def polar(self) -> Energies:
        e = Energies()
        self._check(self.lib.mpmc_b200_polar(self._ctx, C.byref(e)))
        return e

    def mc_steps(self, moves, temperature: float, current_energy: float):
        """Resident MC steps (include/mpmc_b200.h: mpmc_b200_mc_steps): `moves` is a ctypes array of McMove; returns the array
        of McResult (energies of every trial configuration, the device's decisions, trial / old coordinates)."""
        n = len(moves)
        res = (McResult * n)()
        self._check(self.lib.mpmc_b200_mc_steps(self._ctx, n, moves, C.c_double(temperature), C.c_double(current_energy), res))
        return res

    def energy_enqueue(self):
        self._check(self.lib.mpmc_b200_energy_enqueue(self._ctx))

    def energy_finish(self) -> Energies:
        e = Energies()
        self._check(self.lib.mpmc_b200_energy_finish(self._ctx, C.byref(e)))
        return e

    def polar_stage(self, stage: int, iteration: int = 0):
        self._check(self.lib.mpmc_b200_polar_stage(self._ctx, stage, iteration))

    def uses_symmetric_contraction(self) -> bool:
        return bool(self.lib.mpmc_b200_uses_symmetric_contraction(self._ctx))

    def atom_arrays_complete(self) -> bool:
        return bool(self.lib.mpmc_b200_atom_arrays_complete(self._ctx))

    def peer_export(self) -> bytes:
        buf = C.create_string_buffer(64)
        self._check(self.lib.mpmc_b200_peer_export(self._ctx, buf))
        return buf.raw

    def peer_connect(self, handles) -> None:
        blob = b"".join(handles)
        self._check(self.lib.mpmc_b200_peer_connect(self._ctx, len(handles), blob))

    def peer_active(self) -> bool:
        return bool(self.lib.mpmc_b200_peer_active(self._ctx))

    def uses_cluster_field(self) -> bool:
        return bool(self.lib.mpmc_b200_uses_cluster_field(self._ctx))

    def device_buffer(self, which: int):
        ptr, n = C.c_void_p(), C.c_longlong()
        self._check(self.lib.mpmc_b200_device_buffer(self._ctx, which, C.byref(ptr), C.byref(n)))
        return ptr.value, n.value

    def shard_rows(self):
        b, e, n, npad = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        self._check(self.lib.mpmc_b200_shard_rows(self._ctx, C.byref(b), C.byref(e), C.byref(n), C.byref(npad)))
        return b.value, e.value, n.value, npad.value

    def shard_field_rows(self):
        b, e, npad = C.c_int(), C.c_int(), C.c_int()
        self._check(self.lib.mpmc_b200_shard_field_rows(self._ctx, C.byref(b), C.byref(e), C.byref(npad)))
        return b.value, e.value, npad.value

    def atom_arrays(self, out: Optional[dict] = None) -> dict:
        """Per-atom results in reference list order.  `out` (a dict returned by an earlier call for the
        same atom count) is refilled in place, like the write-back buffers a C caller keeps."""
        n = self.n
        if out is None or out["dipole_rrms"].shape[0] != n:
            out = {k: np.empty((n, 3)) for k in ("mu", "ef_static", "ef_induced", "ef_induced_change")}
            out["dipole_rrms"] = np.empty(n)
            out["rank_metric"] = np.empty(n)
        self._check(self.lib.mpmc_b200_get_atom_arrays(self._ctx, _dp(out["mu"]), _dp(out["ef_static"]),
                                                       _dp(out["ef_induced"]), _dp(out["ef_induced_change"]),
                                                       _dp(out["dipole_rrms"]), _dp(out["rank_metric"])))
        return out

    def counters(self) -> Counters:
        c = Counters()
        self._check(self.lib.mpmc_b200_get_counters(self._ctx, C.byref(c)))
        return c

    def phase_ms(self, phase: str) -> float:
        v = C.c_double()
        self._check(self.lib.mpmc_b200_phase_ms(self._ctx, phase.encode(), C.byref(v)))
        return v.value

    def measure_fp64_peak(self) -> float:
        v = C.c_double()
        self._check(self.lib.mpmc_b200_measure_fp64_peak(self._ctx, C.byref(v)))
        return v.value
