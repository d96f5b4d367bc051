#!/usr/bin/env python
"""oracle/make_golden.py -- TEST INFRASTRUCTURE ONLY.

Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref, built by
oracle/Makefile from /root/reference) so the fixtures can travel to the GPU box,
where /root/reference does not exist.  Run in the build container:

    python oracle/make_golden.py            # all cases
    python oracle/make_golden.py mixed_gs   # one case

Each fixture holds the flat System (every input the hot path reads, bit-exact as
the reference parsed it) and everything energy() produced: the observables, the
three Ewald pieces, and per-atom mu / ef_static / ef_induced / ef_induced_change /
dipole_rrms / rank_metric, all at full double precision (ref_driver.c prints %.17g).

Trajectory fixtures (tests/golden/traj_*.json) hold the reference program's own
energy_output rows of a fixed-seed (`preset_seeds`) MC run with `corrtime 1`.
"""
from __future__ import annotations

import json
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mpmc_b200 import systems  # noqa: E402
from mpmc_b200.api import DAMP_LINEAR, DAMP_OFF, MIX_C6, MIX_HALGREN, MIX_WH  # noqa: E402
from oracle import oracle  # noqa: E402

REF = "/root/reference"
GOLD = os.path.join(ROOT, "tests", "golden")


def _variant(base, **kw):
    s = base.copy()
    for k, v in kw.items():
        setattr(s.params, k, v)
    return s


def synthetic_cases():
    m = systems.mixed_small()
    t = systems.mixed_small(triclinic=True)
    prod = dict(polar_gs_ranked=1, polar_max_iter=4, polar_palmo=1, polar_wolf=1, polar_wolf_alpha=0.13)
    c = {
        "mixed_default": m,
        "mixed_tri_default": t,
        "mixed_gs": _variant(m, polar_gs=1),
        "mixed_gs_ranked": _variant(m, polar_gs_ranked=1, polar_max_iter=4),
        "mixed_gs_ranked_palmo_wolf": _variant(m, **prod),
        "mixed_tri_gs_ranked_palmo_wolf": _variant(t, **prod),
        "mixed_palmo": _variant(m, polar_palmo=1),
        "mixed_sor": _variant(m, polar_sor=1, polar_gamma=0.8),
        "mixed_esor": _variant(m, polar_esor=1, polar_gamma=1.2),
        "mixed_gamma": _variant(m, polar_gamma=1.03),
        "mixed_precision": _variant(m, polar_precision=1e-5, polar_max_iter=0),
        "mixed_precision_rrms_gs": _variant(m, polar_precision=1e-6, polar_max_iter=0, polar_rrms=1, polar_gs=1),
        "mixed_rrms": _variant(m, polar_rrms=1, polar_max_iter=3),
        "mixed_zodid": _variant(m, polar_zodid=1),
        "mixed_wolf_a0": _variant(m, polar_wolf=1, polar_wolf_alpha=0.0),
        "mixed_wolf_a013": _variant(m, polar_wolf=1, polar_wolf_alpha=0.13),
        "mixed_wolf_lookup": _variant(m, polar_wolf=1, polar_wolf_alpha=0.13, polar_wolf_alpha_lookup=1),
        "mixed_wolf_full": _variant(m, polar_wolf_full=1, polar_wolf_alpha=0.13),
        "mixed_polar_ewald": _variant(m, polar_ewald=1),
        "mixed_tri_polar_ewald": _variant(t, polar_ewald=1),
        # full-Ewald polarization (ewald_full(): polar_ewald.c:375-418)
        "mixed_ewald_full": _variant(m, polar_ewald_full=1),
        "mixed_tri_ewald_full": _variant(t, polar_ewald_full=1, polar_max_iter=6),
        "mixed_ewald_full_precision_palmo_rrms": _variant(m, polar_ewald_full=1, polar_precision=1e-5, polar_max_iter=0,
                                                          polar_palmo=1, polar_rrms=1),
        "mixed_ewald_full_sor": _variant(m, polar_ewald_full=1, polar_sor=1, polar_gamma=0.8, polar_max_iter=7),
        "mixed_ewald_full_esor_cutoff7": _variant(m, polar_ewald_full=1, polar_esor=1, polar_gamma=1.2, cutoff=7.0,
                                                  ewald_alpha=0.45, polar_ewald_alpha=0.45, polar_palmo=1),
        "h2_fluid_500_ewald_full": _variant(systems.h2_fluid(500), polar_ewald_full=1),
        "mixed_damp_off": _variant(m, damp_type=DAMP_OFF, polar_max_iter=3),
        "mixed_damp_linear": _variant(m, damp_type=DAMP_LINEAR, polar_damp=2.1304),
        "mixed_wh": _variant(m, mixing=MIX_WH),
        "mixed_halgren": _variant(m, mixing=MIX_HALGREN),
        "mixed_c6": _variant(m, mixing=MIX_C6),
        "mixed_fh2": _variant(m, feynman_hibbs_order=2),
        "mixed_nofh_nolrc": _variant(m, feynman_hibbs=0, rd_lrc=0),
        "mixed_nopolar_kmax5": _variant(m, polarization=0, ewald_kmax=5),
        "mixed_rd_only": _variant(m, rd_only=1, polarization=0),
        # Wolf-summed electrostatics (coulombic_wolf(): coulombic.c:269-308); FH must be off (:293-297)
        "mixed_es_wolf": _variant(m, es_wolf=1, feynman_hibbs=0),
        "mixed_tri_es_wolf_nopolar": _variant(t, es_wolf=1, feynman_hibbs=0, polarization=0),
        "co2_fluid_375_es_wolf": _variant(systems.co2_fluid(125), es_wolf=1, feynman_hibbs=0),
        # LJ image sums (rd_crystal: lj.c:109-162,175-216,264-267)
        "mixed_rd_crystal2": _variant(m, rd_crystal=1, rd_crystal_order=2),
        "mixed_tri_rd_crystal2_fh2": _variant(t, rd_crystal=1, rd_crystal_order=2, feynman_hibbs_order=2),
        "mixed_rd_crystal3_nofh_nopolar": _variant(m, rd_crystal=1, rd_crystal_order=3, feynman_hibbs=0, polarization=0),
        "mixed_rd_crystal1_wh": _variant(m, rd_crystal=1, rd_crystal_order=1, mixing=MIX_WH, polarization=0),
        "co2_fluid_375_rd_crystal2": _variant(systems.co2_fluid(125), rd_crystal=1, rd_crystal_order=2),
        "mixed_autoreject_fires": _variant(m, cavity_autoreject_absolute=1, cavity_autoreject_scale=3.2),
        "mixed_autoreject_quiet": _variant(m, cavity_autoreject_absolute=1, cavity_autoreject_scale=0.5),
        "mixed_cutoff7": _variant(m, cutoff=7.0, ewald_alpha=0.45, polar_ewald_alpha=0.45),
        "h2_fluid_500": systems.h2_fluid(500),
        "co2_fluid_375": systems.co2_fluid(125),
        "h2_fluid_500_precision": _variant(systems.h2_fluid(500), polar_precision=1e-5, polar_max_iter=0),
        "h2_fluid_500_precision_palmo_rrms": _variant(systems.h2_fluid(500), polar_precision=1e-7, polar_max_iter=0,
                                                      polar_palmo=1, polar_rrms=1),
    }
    return c


SHIPPED = {
    # name: (directory under the reference, input file, files to copy)
    "ref_small": ("sample_configs_gpu/cuda_pol.small/noncuda_control", "input", ["small.initial.pdb"]),
    "ref_socmof": ("sample_configs_gpu/cuda_pol/noncuda_control", "input", ["socMOF+BSSP.initial.pdb"]),
    "ref_test_lenard_jones": ("mpmc_testing/inputs", "lenard_jones.inp", ["lenard_jones.pqr"]),
    "ref_test_coulombic": ("mpmc_testing/inputs", "coulombic.inp", ["coulombic.pqr"]),
    "ref_test_zero_energy": ("mpmc_testing/inputs", "zero_energy.inp", ["zero_energy.pqr"]),
    "ref_test_infinite_energy": ("mpmc_testing/inputs", "infinite_energy.inp", ["infinite_energy.pqr"]),
}
# option variants of the shipped 1,236-atom system (extra keyword lines appended to its input)
SHIPPED_VARIANTS = {
    "ref_socmof_production": ("ref_socmof", "polar_gs_ranked on\npolar_max_iter 4\npolar_palmo on\npolar_wolf on\n"
                              "polar_wolf_alpha 0.13\n"),
}


def save(name, sysm, exp, extra_meta=None):
    os.makedirs(GOLD, exist_ok=True)
    arrays = sysm.save_arrays()
    scal = {k: v for k, v in exp.items() if not isinstance(v, np.ndarray)}
    arrays["expected_json"] = np.array(json.dumps(scal))
    for k, v in exp.items():
        if isinstance(v, np.ndarray):
            arrays["exp_" + k] = v
    if extra_meta:
        arrays["meta_json"] = np.array(json.dumps(extra_meta))
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), **arrays)


def gen_shipped(name, extra=""):
    base = name if name in SHIPPED else SHIPPED_VARIANTS[name][0]
    d, inp, files = SHIPPED[base]
    with tempfile.TemporaryDirectory() as td:
        for f in files + [inp]:
            shutil.copy(os.path.join(REF, d, f), td)
        if extra:
            with open(os.path.join(td, inp), "a") as f:
                f.write("\n" + extra)
        sysm, exp, _ = oracle.run_reference(os.path.join(td, inp))
    save(name, sysm, exp, {"source": f"{d}/{inp}", "extra": extra})
    return sysm, exp


def gen_synth(name, s):
    with tempfile.TemporaryDirectory() as td:
        inp = systems.write_reference_case(s, td)
        sysm, exp, _ = oracle.run_reference(inp)
    assert np.array_equal(sysm.pos, s.pos) and np.array_equal(sysm.charge, s.charge), name
    save(name, sysm, exp, {"source": "mpmc_b200.systems", "case": name})
    return sysm, exp


TRAJ = {
    # name: (shipped case, extra keywords)
    "traj_small_nvt": ("ref_small", "preset_seeds 12345\nnumsteps 200\ncorrtime 1\n"),
    "traj_small_uvt": ("ref_small", "preset_seeds 777\nnumsteps 300\ncorrtime 1\nensemble uvt\npressure 50.0\n"),
}


def gen_traj(name):
    base, extra = TRAJ[name]
    d, inp, files = SHIPPED[base]
    with tempfile.TemporaryDirectory() as td:
        for f in files + [inp]:
            shutil.copy(os.path.join(REF, d, f), td)
        with open(os.path.join(td, inp), "a") as f:
            f.write("\n" + extra)
        r = subprocess.run([oracle.REF_BINARY, inp], cwd=td, capture_output=True, text=True, timeout=3600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        en = [l for l in open(os.path.join(td, "small.energy.dat")).read().splitlines() if not l.startswith("#")]
        final = open(os.path.join(td, "small.final.pdb")).read()
        inp_text = open(os.path.join(td, inp)).read()
        pdb = open(os.path.join(td, files[0])).read()
    with open(os.path.join(GOLD, name + ".json"), "w") as f:
        json.dump({"input": inp_text, "pqr_name": files[0], "pqr": pdb, "energy_rows": en, "final_pqr": final}, f,
                  indent=0)
    print(f"{name}: {len(en)} energy rows")


# BASELINE config 3: the shipped PCN-61 + H2 system (sample_configs_gpu/3_PCN61: 6,048 frozen framework atoms + 3,027
# five-site H2 = 21,183 atoms, options of iter.inp).  The reference needs ~80 GB for it (pair lists + A matrix), so:
#   ref_pcn61_h2_500   the framework + the first 500 H2 (8,548 atoms) through the UNMODIFIED reference;
#   oracle_pcn61_full  the whole file, parsed here (parser checked against the reference's own parse of the part),
#                      options exactly as the reference parsed iter.inp, expected values from the ORACLE -- the one
#                      fixture not generated by the reference (its name says so; tests treat it as oracle-pinned).
PCN61 = "sample_configs_gpu/3_PCN61"


def parse_pqr(path):
    """Columns of read_molecules() (reference src/io/read_pqr.c:212-218): ATOM id type moltype F|M molid x y z mass q alpha
    eps sig.  Molecules are runs of equal molid, in file order (list order); molecule mass = sum of its atoms' masses."""
    rows = [l.split() for l in open(path) if l.startswith("ATOM")]
    n = len(rows)
    pos = np.array([[float(t[6]), float(t[7]), float(t[8])] for t in rows])
    mass = np.array([float(t[9]) for t in rows])
    q = np.array([float(t[10]) for t in rows]) * systems.E2REDUCED
    alpha = np.array([float(t[11]) for t in rows])
    eps = np.array([float(t[12]) for t in rows])
    sig = np.array([float(t[13]) for t in rows])
    frozen = np.array([1 if t[4] == "F" else 0 for t in rows], dtype=np.int32)
    molid = np.array([int(t[5]) for t in rows])
    mol = np.zeros(n, dtype=np.int32)
    mol[1:] = np.cumsum(molid[1:] != molid[:-1])
    mol_mass = np.zeros(n)
    for m in range(int(mol[-1]) + 1):
        sel = mol == m
        mm = 0.0
        for v in mass[sel]:  # the reference adds the atoms' masses in list order
            mm += v
        mol_mass[sel] = mm
    return pos, q, alpha, eps, sig, mol_mass, mol, frozen


def gen_pcn61(want):
    from mpmc_b200.api import System

    names = ("ref_pcn61_h2_500", "oracle_pcn61_full")
    if want and not (set(names) & want):
        return
    src = os.path.join(REF, PCN61)
    lines = [l for l in open(os.path.join(src, "input.pdb")).read().splitlines() if l.startswith("ATOM")]
    keep, mols = [], []
    for l in lines:
        t = l.split()
        if t[4] != "F":
            if t[5] not in mols:
                if len(mols) >= 500:
                    continue
                mols.append(t[5])
        keep.append(l)
    with tempfile.TemporaryDirectory() as td:
        with open(os.path.join(td, "input.pdb"), "w") as f:
            f.write("\n".join(keep) + "\nEND\n")
        shutil.copy(os.path.join(src, "iter.inp"), td)
        part, exp, _ = oracle.run_reference(os.path.join(td, "iter.inp"))
        mine = parse_pqr(os.path.join(td, "input.pdb"))
    for a, b in zip(mine, (part.pos, part.charge, part.polarizability, part.epsilon, part.sigma, part.mol_mass, part.mol_id,
                           part.frozen)):
        assert np.array_equal(np.asarray(a), np.asarray(b)), "parse_pqr disagrees with the reference's parser"
    if not want or "ref_pcn61_h2_500" in want:
        save("ref_pcn61_h2_500", part, exp, {"source": f"{PCN61}/iter.inp", "extra": "framework + the first 500 H2 of input.pdb"})
        print(f"ref_pcn61_h2_500: n={part.n} total={exp['total']!r} polar={exp['polar']!r}")
    if not want or "oracle_pcn61_full" in want:
        pos, q, alpha, eps, sig, mm, mol, frozen = parse_pqr(os.path.join(src, "input.pdb"))
        full = System(part.params, pos, q, alpha, eps, sig, mm, mol, frozen, "oracle_pcn61_full")
        e, _, arr = oracle.energy(full, nthreads=0)
        expf = e.to_dict()
        expf.update({k: arr[k] for k in ("mu", "ef_static", "ef_induced", "ef_induced_change", "dipole_rrms", "rank_metric")})
        expf["polar_ewald_full"] = 0
        save("oracle_pcn61_full", full, expf, {"source": f"{PCN61}/input.pdb + iter.inp", "pinned_by": "oracle (oracle/mpmc_oracle.c); "
                                              "the reference needs ~80 GB for this system"})
        print(f"oracle_pcn61_full: n={full.n} total={expf['total']!r} polar={expf['polar']!r} iters={expf['polar_iterations']}")


# fixed-seed trajectories of synthetic systems: name: (builder, ensemble, extra keywords)
TRAJ_SYNTH = {
    # LJ + Ewald, no polarization (BASELINE config 2 in small): the case the incremental path applies to
    "traj_co2_nvt": (lambda: systems.co2_fluid(125), "nvt", "preset_seeds 4242\nnumsteps 300\nmove_factor 0.05\nrot_factor 0.05"),
    # the same with insertions / removals (N goes 125 -> 99): every change of the atom set must drop the incremental state
    "traj_co2_uvt": (lambda: systems.co2_fluid(125), "uvt", "preset_seeds 99\nnumsteps 300\nmove_factor 0.05\nrot_factor 0.05\n"
                                                            "pressure 200.0\ninsert_probability 0.4"),
    # volume moves (V goes 8000 -> 9006 A^3): the cutoff, alpha, k-vectors and LRC follow the cell; the input leaves the
    # cutoff to the reference's pbc() so that it is recomputed with every volume change
    "traj_co2_npt": (lambda: systems.co2_fluid(125), "npt", "preset_seeds 31337\nnumsteps 200\nmove_factor 0.05\nrot_factor 0.05\n"
                                                            "pressure 100.0\nvolume_probability 0.2\nvolume_change_factor 0.02",
                     dict(explicit_cutoff=False)),
    # the remaining relink shapes of the MC driver: cavity-biased insertions (mc_moves.c:577-600, cavity.c) and
    # insertions of molecules taken from a separate list (`insert_input`: the new molecule is a copy of a template that
    # was never in the system, mc_moves.c:603-660)
    "traj_co2_uvt_cavity": (lambda: systems.co2_fluid(125), "uvt", "preset_seeds 2718\nnumsteps 300\nmove_factor 0.05\nrot_factor 0.05\n"
                                                                   "pressure 200.0\ninsert_probability 0.4\ncavity_bias on\n"
                                                                   "cavity_grid 5\ncavity_radius 3.0"),
    "traj_co2_uvt_insert_input": (lambda: systems.co2_fluid(125), "uvt", "preset_seeds 1618\nnumsteps 300\nmove_factor 0.05\n"
                                                                         "rot_factor 0.05\npressure 200.0\ninsert_probability 0.4\n"
                                                                         "insert_input insert.pqr", dict(insert_first_molecule=True)),
}


def gen_traj_synth(name):
    build, ensemble, extra = TRAJ_SYNTH[name][:3]
    kw = dict(TRAJ_SYNTH[name][3]) if len(TRAJ_SYNTH[name]) > 3 else {}
    insert_first = kw.pop("insert_first_molecule", False)
    s = build()
    first = np.r_[True, s.mol_id[1:] != s.mol_id[:-1]]
    s.meta["atom_mass"] = np.where(first, s.mol_mass, 0.0)  # update_com() then recovers molecule_t.mass
    with tempfile.TemporaryDirectory() as td:
        inp = systems.write_reference_case(s, td, ensemble=ensemble, extra=extra, **kw)
        extra_files = {}
        if insert_first:  # the template list of `insert_input`: the first molecule of the system, on its own
            lines = [l for l in open(os.path.join(td, "system.pqr")).read().splitlines() if l.startswith("ATOM")]
            first = [l for l in lines if l.split()[5] == lines[0].split()[5]]
            extra_files["insert.pqr"] = "\n".join(first) + "\nEND\n"
        for fn, text in extra_files.items():
            with open(os.path.join(td, fn), "w") as f:
                f.write(text)
        r = subprocess.run([oracle.REF_BINARY, "input"], cwd=td, capture_output=True, text=True, timeout=3600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        en = [l for l in open(os.path.join(td, "energy.dat")).read().splitlines() if not l.startswith("#")]
        final = open(os.path.join(td, "final.pqr")).read()
        inp_text = open(inp).read()
        pqr = open(os.path.join(td, "system.pqr")).read()
    with open(os.path.join(GOLD, name + ".json"), "w") as f:
        json.dump({"input": inp_text, "pqr_name": "system.pqr", "pqr": pqr, "energy_rows": en, "final_pqr": final,
                   "energy_file": "energy.dat", "final_file": "final.pqr", "extra_files": extra_files}, f, indent=0)
    print(f"{name}: {len(en)} energy rows")


# n2 (north_star: "ensemble averages and uptake of the sample MOF runs must agree"): the shipped PCN-61 + H2 uVT input
# (sample_configs_gpu/3_PCN61/iter.inp, all of its options) on the largest part of the system the reference can hold --
# the framework + the first 500 H2 (8,548 atoms, ~13 GB) -- for 100 fixed-seed steps: every energy row AND the
# `OUTPUT:` averages block the reference prints at each corrtime (N, wt %, excess ratio, <U>, <U_pol>, qst, AR ...).
def gen_traj_pcn61():
    src = os.path.join(REF, PCN61)
    lines = [l for l in open(os.path.join(src, "input.pdb")).read().splitlines() if l.startswith("ATOM")]
    keep, mols = [], []
    for l in lines:
        t = l.split()
        if t[4] != "F":
            if t[5] not in mols:
                if len(mols) >= 500:
                    continue
                mols.append(t[5])
        keep.append(l)
    inp = open(os.path.join(src, "iter.inp")).read()
    inp = inp.replace("numsteps\t\t4", "numsteps\t\t100").replace("corrtime\t\t2", "corrtime\t\t10")
    assert "numsteps\t\t100" in inp and "corrtime\t\t10" in inp
    inp += "\npreset_seeds 20261105\n"
    with tempfile.TemporaryDirectory() as td:
        with open(os.path.join(td, "input.pdb"), "w") as f:
            f.write("\n".join(keep) + "\nEND\n")
        with open(os.path.join(td, "input"), "w") as f:
            f.write(inp)
        r = subprocess.run([oracle.REF_BINARY, "input"], cwd=td, capture_output=True, text=True, timeout=6 * 3600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        en = [l for l in open(os.path.join(td, "RHT+BSSP.energy.dat")).read().splitlines() if not l.startswith("#")]
        final = open(os.path.join(td, "RHT+BSSP.final.pdb")).read()
        dip = open(os.path.join(td, "RHT+BSSP.dipole.dat")).read() if os.path.exists(os.path.join(td, "RHT+BSSP.dipole.dat")) else ""
    out_lines = [l for l in r.stdout.splitlines() if l.startswith("OUTPUT:") and "sec/step" not in l and "collecting statistics" not in l]
    import gzip
    import base64

    pack = lambda t: base64.b64encode(gzip.compress(t.encode(), 9)).decode()
    with open(os.path.join(GOLD, "traj_pcn61_uvt.json"), "w") as f:
        json.dump({"input": inp, "pqr_name": "input.pdb", "pqr_gz_b64": pack("\n".join(keep) + "\nEND\n"), "energy_rows": en,
                   "final_gz_b64": pack(final), "dipole_tail": dip.splitlines()[-3:], "output_lines": out_lines,
                   "energy_file": "RHT+BSSP.energy.dat", "final_file": "RHT+BSSP.final.pdb",
                   "source": f"{PCN61}/iter.inp + framework and the first 500 H2 of input.pdb; numsteps 100, corrtime 10, "
                             "preset_seeds 20261105; unmodified reference (oracle/_ref/mpmc_ref)"}, f, indent=0)
    print(f"traj_pcn61_uvt: {len(en)} energy rows, {len(out_lines)} OUTPUT lines")


def main():
    oracle.build()
    want = set(sys.argv[1:])
    for name in list(SHIPPED) + list(SHIPPED_VARIANTS):
        if want and name not in want:
            continue
        extra = SHIPPED_VARIANTS[name][1] if name in SHIPPED_VARIANTS else ""
        try:
            s, e = gen_shipped(name, extra)
            print(f"{name}: n={s.n} total={e['total']!r}")
        except Exception as ex:  # noqa: BLE001
            print(f"{name}: FAILED {str(ex)[-600:]}")
    for name, s in synthetic_cases().items():
        if want and name not in want:
            continue
        try:
            sr, e = gen_synth(name, s)
            print(f"{name}: n={sr.n} total={e['total']!r} polar={e['polar']!r} iters={e['polar_iterations']}")
        except Exception as ex:  # noqa: BLE001
            print(f"{name}: FAILED {str(ex)[-600:]}")
    gen_pcn61(want)
    for name in TRAJ:
        if want and name not in want:
            continue
        gen_traj(name)
    for name in TRAJ_SYNTH:
        if want and name not in want:
            continue
        gen_traj_synth(name)
    if "traj_pcn61_uvt" in want:  # ~30 minutes of the reference on one core: only on request
        gen_traj_pcn61()


if __name__ == "__main__":
    main()
