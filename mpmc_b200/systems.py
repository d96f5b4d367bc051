"""Synthetic systems of BASELINE.json's configs and
writers for the reference's own file formats (PQR + keyword input), so that the
unmodified reference can be run on exactly the same system.

All generators are seeded ``numpy.random.Generator(PCG64(20261104 + k))``.
Nothing here reads /root/reference.
"""
from __future__ import annotations

import os
from typing import Optional

import numpy as np

from .api import DAMP_EXPONENTIAL, E2REDUCED, Params, System, default_params

SEED = 20261104

# H2 "BSSP" five-site model -- parameter values as in the reference's sample
# sample_configs_gpu/cuda_pol.small/small.initial.pdb:1-5 (site, z offset along
# the bond in A, mass, charge/e, alpha, epsilon, sigma)
H2_BSSP = [
    ("H2G", 0.000, 0.0000, -0.7464, 0.69380, 12.76532, 3.15528),
    ("H2E", +0.371, 1.0080, 0.3732, 0.00044, 0.00000, 0.00000),
    ("H2E", -0.371, 1.0080, 0.3732, 0.00044, 0.00000, 0.00000),
    ("H2N", +0.363, 0.0000, 0.0000, 0.00000, 2.16726, 2.37031),
    ("H2N", -0.363, 0.0000, 0.0000, 0.00000, 2.16726, 2.37031),
]
# CO2-like rigid 3-site molecule of config M2
CO2_LIKE = [
    ("C", 0.00, 12.011, +0.70, 0.0, 27.0, 2.80),
    ("O", +1.16, 15.999, -0.35, 0.0, 79.0, 3.05),
    ("O", -1.16, 15.999, -0.35, 0.0, 79.0, 3.05),
]
# framework "elements" (name, mass, alpha, epsilon, sigma): UFF-like LJ values; polarizabilities are
# the usual C/O/H/Cu point polarizabilities (1.2886, 0.852, 0.4138, 2.1963 A^3) halved, which makes the
# synthetic scaffold's Thole iteration matrix contractive (spectral radius ~0.73, so the Jacobi solve
# of the reference converges; with the full values on this artificial geometry it diverges).
FRAMEWORK_TYPES = [
    ("C", 12.011, 0.64430, 52.84, 3.431),
    ("O", 15.999, 0.42600, 30.19, 3.118),
    ("H", 1.008, 0.20690, 22.14, 2.571),
    ("CU", 63.546, 1.09815, 2.516, 3.114),
]


def _random_unit_vectors(rng, m):
    v = rng.normal(size=(m, 3))
    return v / np.linalg.norm(v, axis=1, keepdims=True)


def _molecules(centers, axes, model, mol_id0=0, frozen=0):
    """Place a rigid linear model on each centre along each axis."""
    m = centers.shape[0]
    ns = len(model)
    off = np.array([s[1] for s in model])
    pos = (centers[:, None, :] + off[None, :, None] * axes[:, None, :]).reshape(m * ns, 3)
    mass = np.tile([s[2] for s in model], m)
    q = np.tile([s[3] for s in model], m) * E2REDUCED
    alpha = np.tile([s[4] for s in model], m)
    eps = np.tile([s[5] for s in model], m)
    sig = np.tile([s[6] for s in model], m)
    mol = np.repeat(np.arange(m, dtype=np.int32) + mol_id0, ns)
    types = [s[0] for s in model] * m
    molmass = np.full(m * ns, float(sum(s[2] for s in model)))
    return dict(pos=pos, mass=mass, q=q, alpha=alpha, eps=eps, sig=sig, mol=mol, types=types, molmass=molmass,
                frozen=np.full(m * ns, frozen, dtype=np.int32), moltype=["H2" if model is H2_BSSP else "CO2"] * (m * ns))


def _assemble(parts, params, name) -> System:
    cat = lambda k: np.concatenate([p[k] for p in parts])
    s = System(params, cat("pos"), cat("q"), cat("alpha"), cat("eps"), cat("sig"), cat("molmass"), cat("mol"),
               cat("frozen"), name)
    s.meta["atom_mass"] = cat("mass")
    s.meta["atomtype"] = sum((p["types"] for p in parts), [])
    s.meta["moltype"] = sum((p["moltype"] for p in parts), [])
    return s


def _lattice_centres(rng, n_mol, a, jitter):
    g = int(np.ceil(n_mol ** (1.0 / 3.0)))
    idx = np.stack(np.meshgrid(np.arange(g), np.arange(g), np.arange(g), indexing="ij"), -1).reshape(-1, 3)[:n_mol]
    L = g * a
    c = (idx + 0.5) * a - 0.5 * L + rng.uniform(-jitter, jitter, size=(n_mol, 3))
    return c, L


def h2_fluid(n_atoms: int, seed: int = SEED, polar_max_iter: int = 10) -> System:
    """M5: all-mobile H2-BSSP fluid, N/5 molecules on a jittered simple-cubic lattice a=3.6 A,
    options of cuda_pol.small (FH4, LRC, exponential damping 2.1304, Jacobi x10, bare-Coulomb field)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    m = n_atoms // 5
    c, L = _lattice_centres(rng, m, 3.6, 0.3)
    part = _molecules(c, _random_unit_vectors(rng, m), H2_BSSP)
    p = default_params(np.eye(3) * L)
    p.temperature = 77.0
    p.feynman_hibbs, p.feynman_hibbs_order = 1, 4
    p.polarization, p.polar_iterative = 1, 1
    p.damp_type, p.polar_damp = DAMP_EXPONENTIAL, 2.1304
    p.polar_max_iter = polar_max_iter
    return _assemble([part], p, f"h2_fluid_{m * 5}")


def co2_fluid(n_mol: int = 4096, seed: int = SEED + 1) -> System:
    """M2: LJ+Ewald fluid, no polarization: rigid 3-site molecules on a 16^3 lattice (a=4.0 A)
    with +-0.3 A jitter, cubic box, default cutoff L/2, alpha 3.5/cutoff, kmax 7, rd_lrc on, T=300 K."""
    rng = np.random.Generator(np.random.PCG64(seed))
    c, L = _lattice_centres(rng, n_mol, 4.0, 0.3)
    part = _molecules(c, _random_unit_vectors(rng, n_mol), CO2_LIKE)
    p = default_params(np.eye(3) * L)
    p.temperature = 300.0
    return _assemble([part], p, f"co2_fluid_{n_mol * 3}")


def mof_h2(cells=(9, 9, 3), cell_edge: float = 21.398 * 2.0 / 3.0, n_h2: Optional[int] = None, seed: int = SEED + 2,
           cutoff: float = 8.0, polar_max_iter: int = 10, polar_precision: float = 0.0) -> System:
    """M4-like synthetic MOF supercell + H2: a frozen, neutral, polarizable open scaffold (one frozen
    molecule, listed first, as MOF PQR files do) filled with H2-BSSP molecules.  Defaults give
    N ~ 63k atoms in a 128.388 x 128.388 x 42.796 A box (~18k frozen atoms, 9,072 H2 => ~45k
    polarizable sites), options of sample_configs_gpu/3_PCN61/iter.inp (FH4, exponential damping
    2.1304, pbc_cutoff 8.0, Jacobi, bare-Coulomb static field)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    cells = np.asarray(cells)
    box = cells * cell_edge
    # --- scaffold: sites of a simple-cubic grid (spacing = cell_edge/10 ~ 1.43 A, a bonded distance).
    # Every cell edge carries a hollow rod of three single-file chains, 2.85 A apart, so an atom has
    # two bonded neighbours (more only where rods cross) -- a dense solid rod makes the Thole
    # iteration matrix expansive (Jacobi diverges), a chain-like one does not.  Sites are jittered
    # by < 0.1 A.
    m = 10
    ag = cell_edge / m
    gi = np.stack(np.meshgrid(np.arange(cells[0] * m), np.arange(cells[1] * m), np.arange(cells[2] * m),
                              indexing="ij"), -1).reshape(-1, 3)
    loc = gi % m
    off = np.where(loc > m // 2, loc - m, loc)  # signed offset from the nearest lower/upper cell face
    keep = np.zeros(gi.shape[0], dtype=bool)
    for ax in range(3):
        o1, o2 = off[:, (ax + 1) % 3], off[:, (ax + 2) % 3]
        keep |= ((np.abs(o1) == 1) & (np.abs(o2) == 1)) & ~((o1 == -1) & (o2 == -1))  # 3 of the 4 corner chains
    gi = gi[keep]
    h = (gi[:, 0] * 73856093) ^ (gi[:, 1] * 19349663) ^ (gi[:, 2] * 83492791)
    kind = np.select([h % 20 == 0, h % 20 < 5, h % 20 < 12], [3, 1, 2], default=0)  # 5% Cu, 20% O, 35% H, 40% C
    pts = [gi * ag + rng.uniform(-0.07, 0.07, size=gi.shape)]
    kinds = [kind]
    fpos = np.concatenate(pts)
    fkind = np.concatenate(kinds).astype(int)
    fpos = fpos - 0.5 * box  # centre the box on the origin
    nf = fpos.shape[0]
    falpha = np.array([FRAMEWORK_TYPES[k][2] for k in fkind])
    feps = np.array([FRAMEWORK_TYPES[k][3] for k in fkind])
    fsig = np.array([FRAMEWORK_TYPES[k][4] for k in fkind])
    fmass = np.array([FRAMEWORK_TYPES[k][1] for k in fkind])
    fq = np.array([+0.35, -0.55, +0.15, +1.05])[fkind] + rng.uniform(-0.05, 0.05, size=nf)
    fq -= fq.mean()  # neutral framework
    fpart = dict(pos=fpos, mass=fmass, q=fq * E2REDUCED, alpha=falpha, eps=feps, sig=fsig,
                 mol=np.zeros(nf, dtype=np.int32), types=[FRAMEWORK_TYPES[k][0] for k in fkind],
                 molmass=np.full(nf, fmass.sum()), frozen=np.ones(nf, dtype=np.int32), moltype=["MOF"] * nf)
    # --- sorbate: H2 centres at least 3.0 A from any scaffold atom and 2.6 A from each other
    from scipy.spatial import cKDTree

    if n_h2 is None:
        n_h2 = int(round(9072 * np.prod(cells) / 243.0))
    wrap = lambda x: (x + 0.5 * box) % box
    ftree = cKDTree(wrap(fpos), boxsize=box)
    centres = np.zeros((0, 3))
    while centres.shape[0] < n_h2:
        cand = rng.uniform(-0.5, 0.5, size=(4 * n_h2, 3)) * box
        d, _ = ftree.query(wrap(cand), k=1)
        cand = cand[d > 3.0]
        allc = np.concatenate([centres, cand])
        tree = cKDTree(wrap(allc), boxsize=box)
        pairs = tree.query_pairs(2.6, output_type="ndarray")
        drop = np.zeros(allc.shape[0], dtype=bool)
        for i, j in pairs:  # greedy: drop the later one
            if not drop[i] and not drop[j]:
                drop[max(i, j)] = True
        centres = allc[~drop][:n_h2]
    hpart = _molecules(centres, _random_unit_vectors(rng, n_h2), H2_BSSP, mol_id0=1)
    p = default_params(np.diag(box), cutoff=cutoff)
    p.temperature = 77.0
    p.feynman_hibbs, p.feynman_hibbs_order = 1, 4
    p.polarization, p.polar_iterative = 1, 1
    p.damp_type, p.polar_damp = DAMP_EXPONENTIAL, 2.1304
    p.polar_max_iter = polar_max_iter
    p.polar_precision = polar_precision
    return _assemble([fpart, hpart], p, f"mof_h2_{nf + 5 * n_h2}")


def mixed_small(seed: int = SEED + 3, triclinic: bool = False, n_h2: int = 40, n_frozen: int = 60) -> System:
    """Small test system with every kind of atom the path distinguishes: a frozen charged
    polarizable cluster (one molecule), H2-BSSP molecules, CO2-like molecules (non-polarizable),
    in an orthorhombic or triclinic cell."""
    rng = np.random.Generator(np.random.PCG64(seed))
    basis = np.array([[24.0, 0.0, 0.0], [0.0, 22.0, 0.0], [0.0, 0.0, 26.0]])
    if triclinic:
        basis = np.array([[24.0, 0.0, 0.0], [3.0, 22.0, 0.0], [-2.0, 4.0, 26.0]])
    # frozen cluster on a jittered 1.6 A grid near the origin
    g = np.stack(np.meshgrid(np.arange(5), np.arange(4), np.arange(3), indexing="ij"), -1).reshape(-1, 3)[:n_frozen]
    fpos = (g - g.mean(0)) * 1.7 + rng.uniform(-0.15, 0.15, size=(n_frozen, 3))
    kind = rng.integers(0, 4, size=n_frozen)
    fq = rng.uniform(-0.6, 0.6, size=n_frozen)
    fq -= fq.mean()
    fmass = np.array([FRAMEWORK_TYPES[k][1] for k in kind])
    fpart = dict(pos=fpos, mass=fmass, q=fq * E2REDUCED, alpha=np.array([FRAMEWORK_TYPES[k][2] for k in kind]),
                 eps=np.array([FRAMEWORK_TYPES[k][3] for k in kind]),
                 sig=np.array([FRAMEWORK_TYPES[k][4] for k in kind]), mol=np.zeros(n_frozen, dtype=np.int32),
                 types=[FRAMEWORK_TYPES[k][0] for k in kind], molmass=np.full(n_frozen, fmass.sum()),
                 frozen=np.ones(n_frozen, dtype=np.int32), moltype=["MOF"] * n_frozen)
    # sorbates at random fractional coordinates, kept 3.2 A away from the cluster and 2.8 A apart
    centres = []
    while len(centres) < n_h2 + 10:
        f = rng.uniform(-0.5, 0.5, size=3)
        c = f @ basis
        if np.min(np.linalg.norm(fpos - c, axis=1)) < 5.0 and np.linalg.norm(c) < 6.5:
            continue
        ok = True
        for o in centres:
            d = c - o
            fr = np.linalg.solve(basis.T, d)
            d = d - np.rint(fr) @ basis
            if np.linalg.norm(d) < 2.8:
                ok = False
                break
        if ok:
            centres.append(c)
    centres = np.array(centres)
    hpart = _molecules(centres[:n_h2], _random_unit_vectors(rng, n_h2), H2_BSSP, mol_id0=1)
    cpart = _molecules(centres[n_h2:], _random_unit_vectors(rng, centres.shape[0] - n_h2), CO2_LIKE,
                       mol_id0=1 + n_h2)
    p = default_params(basis)
    p.temperature = 77.0
    p.feynman_hibbs, p.feynman_hibbs_order = 1, 4
    p.polarization, p.polar_iterative = 1, 1
    p.damp_type, p.polar_damp = DAMP_EXPONENTIAL, 2.1304
    return _assemble([fpart, hpart, cpart], p, "mixed_small" + ("_tri" if triclinic else ""))


# ---------------------------------------------------------------------------
# writers for the reference's file formats
# ---------------------------------------------------------------------------
def write_pqr(s: System, path: str):
    """PQR in the column order read_molecules() parses (reference src/io/read_pqr.c:212-218):
    ATOM id type moltype F|M molid x y z mass charge(e) alpha epsilon sigma.  %.17g everywhere
    so the reference sees bit-identical inputs."""
    mass = s.meta.get("atom_mass", np.zeros(s.n))
    types = s.meta.get("atomtype", ["X"] * s.n)
    mt = s.meta.get("moltype", ["MOL"] * s.n)
    with open(path, "w") as f:
        for i in range(s.n):
            f.write("ATOM %d %s %s %s %d %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g\n" %
                    (i + 1, types[i], mt[i], "F" if s.frozen[i] else "M", int(s.mol_id[i]) + 1, s.pos[i, 0], s.pos[i, 1],
                     s.pos[i, 2], mass[i], s.charge[i] / E2REDUCED, s.polarizability[i], s.epsilon[i], s.sigma[i]))
        f.write("END\n")


def write_input(s: System, path: str, pqr_name: str, ensemble: str = "total_energy", extra: str = "",
                explicit_cutoff: bool = True):
    """Keyword input file (reference src/io/input.c:73-1596) reproducing s.params."""
    p = s.params
    L = []
    a = L.append
    a(f"job_name {s.name or 'job'}")
    a(f"ensemble {ensemble}")
    a("numsteps 1")
    a("corrtime 1")
    a(f"temperature {p.temperature!r}")
    for q in range(3):
        a("basis%d %.17g %.17g %.17g" % (q + 1, p.basis[q][0], p.basis[q][1], p.basis[q][2]))
    if explicit_cutoff:
        a(f"pbc_cutoff {p.cutoff!r}")
        a(f"ewald_alpha {p.ewald_alpha!r}")
        a(f"polar_ewald_alpha {p.polar_ewald_alpha!r}")
    a(f"ewald_kmax {p.ewald_kmax}")
    a(f"rd_lrc {'on' if p.rd_lrc else 'off'}")
    if p.rd_only:
        a("rd_only on")
    if p.es_wolf:
        a("wolf on")
    if p.rd_crystal:
        a("rd_crystal on")
        a(f"rd_crystal_order {p.rd_crystal_order}")
    if p.feynman_hibbs:
        a("feynman_hibbs on")
        a(f"feynman_hibbs_order {p.feynman_hibbs_order}")
    if p.mixing == 1:
        a("waldmanhagler on")
    elif p.mixing == 2:
        a("halgren_mixing on")
    elif p.mixing == 3:
        a("c6_mixing on")
    if p.cavity_autoreject_absolute:
        a("cavity_autoreject_absolute on")
        a(f"cavity_autoreject_scale {p.cavity_autoreject_scale!r}")
    if p.polarization:
        a("polarization on")
        a("polar_damp_type " + ["off", "linear", "exponential"][p.damp_type])
        a(f"polar_damp {p.polar_damp!r}")
        if p.polar_iterative:
            a("polar_iterative on")
        if p.polar_precision > 0:
            a("polar_max_iter 0")
            a(f"polar_precision {p.polar_precision!r}")
        else:
            a(f"polar_max_iter {p.polar_max_iter}")
        a(f"polar_gamma {p.polar_gamma!r}")
        for k in ("polar_gs", "polar_gs_ranked", "polar_sor", "polar_esor", "polar_palmo", "polar_rrms",
                  "polar_zodid", "polar_ewald", "polar_ewald_full", "polar_wolf", "polar_wolf_full"):
            if getattr(p, k):
                a(f"{k} on")
        if p.polar_wolf or p.polar_wolf_full:
            a(f"polar_wolf_alpha {p.polar_wolf_alpha!r}")
            if p.polar_wolf_alpha_lookup:
                a("polar_wolf_alpha_lookup on")
                a(f"polar_wolf_alpha_lookup_cutoff {p.polar_wolf_alpha_lookup_cutoff!r}")
    a(f"pqr_input {pqr_name}")
    a("pqr_output final.pqr")
    a("pqr_restart restart.pqr")
    a("energy_output energy.dat")
    a("traj_output /dev/null")
    a("wrapall on")
    if extra:
        a(extra)
    with open(path, "w") as f:
        f.write("\n".join(L) + "\n")


def write_reference_case(s: System, workdir: str, **kw) -> str:
    os.makedirs(workdir, exist_ok=True)
    write_pqr(s, os.path.join(workdir, "system.pqr"))
    inp = os.path.join(workdir, "input")
    write_input(s, inp, "system.pqr", **kw)
    return inp
