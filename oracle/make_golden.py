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


# fixed-seed trajectories of synthetic systems: name: (builder, ensemble, extra keywords)
TRAJ_SYNTH = {
    # LJ + Ewald, no polarization (BASELINE config 2 in small): the case the incremental path applies to
    "traj_co2_nvt": (lambda: systems.co2_fluid(125), "nvt", "preset_seeds 4242\nnumsteps 300\nmove_factor 0.05\nrot_factor 0.05"),
    # the same with insertions / removals (N goes 125 -> 99): every change of the atom set must drop the incremental state
    "traj_co2_uvt": (lambda: systems.co2_fluid(125), "uvt", "preset_seeds 99\nnumsteps 300\nmove_factor 0.05\nrot_factor 0.05\n"
                                                            "pressure 200.0\ninsert_probability 0.4"),
}


def gen_traj_synth(name):
    build, ensemble, extra = TRAJ_SYNTH[name]
    s = build()
    first = np.r_[True, s.mol_id[1:] != s.mol_id[:-1]]
    s.meta["atom_mass"] = np.where(first, s.mol_mass, 0.0)  # update_com() then recovers molecule_t.mass
    with tempfile.TemporaryDirectory() as td:
        inp = systems.write_reference_case(s, td, ensemble=ensemble, extra=extra)
        r = subprocess.run([oracle.REF_BINARY, "input"], cwd=td, capture_output=True, text=True, timeout=3600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        en = [l for l in open(os.path.join(td, "energy.dat")).read().splitlines() if not l.startswith("#")]
        final = open(os.path.join(td, "final.pqr")).read()
        inp_text = open(inp).read()
        pqr = open(os.path.join(td, "system.pqr")).read()
    with open(os.path.join(GOLD, name + ".json"), "w") as f:
        json.dump({"input": inp_text, "pqr_name": "system.pqr", "pqr": pqr, "energy_rows": en, "final_pqr": final,
                   "energy_file": "energy.dat", "final_file": "final.pqr"}, f, indent=0)
    print(f"{name}: {len(en)} energy rows")


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
    for name in TRAJ:
        if want and name not in want:
            continue
        gen_traj(name)
    for name in TRAJ_SYNTH:
        if want and name not in want:
            continue
        gen_traj_synth(name)


if __name__ == "__main__":
    main()
