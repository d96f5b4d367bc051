"""CPU-only checks of the boundary: the C-ABI library loads, exports every symbol the header
declares, the ctypes mirrors match the C structs, and it fails loudly without a GPU."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

import mpmc_b200
from mpmc_b200 import api, systems

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(mpmc_b200.LIB_PATH):
        import __graft_entry__ as g

        g.build()
    return ctypes.CDLL(mpmc_b200.LIB_PATH)


def test_exports_every_declared_symbol(lib):
    hdr = open(os.path.join(ROOT, "include", "mpmc_b200.h")).read()
    declared = sorted(set(re.findall(r"\b(mpmc_b200_[a-z0-9_]+)\s*\(", hdr)))
    assert declared, "no prototypes found"
    assert sorted(api.ABI_SYMBOLS) == declared
    for name in declared:
        assert hasattr(lib, name), name


def test_struct_layouts_match_header(tmp_path):
    """sizeof/offsetof of the C structs (compiled with gcc) equal the ctypes mirrors."""
    src = tmp_path / "layout.c"
    fields_p = [n for n, _ in api.Params._fields_]
    fields_e = [n for n, _ in api.Energies._fields_]
    fields_c = [n for n, _ in api.Counters._fields_]
    body = ['#include <stdio.h>', '#include <stddef.h>', '#include "mpmc_b200.h"', 'int main(void){']
    body.append('printf("%zu %zu %zu\\n", sizeof(mpmc_b200_params), sizeof(mpmc_b200_energies), sizeof(mpmc_b200_counters));')
    for f in fields_p:
        body.append(f'printf("%zu\\n", offsetof(mpmc_b200_params, {f}));')
    for f in fields_e:
        body.append(f'printf("%zu\\n", offsetof(mpmc_b200_energies, {f}));')
    for f in fields_c:
        body.append(f'printf("%zu\\n", offsetof(mpmc_b200_counters, {f}));')
    fields_m = [n for n, _ in api.McMove._fields_]
    fields_r = [n for n, _ in api.McResult._fields_]
    body.append('printf("%zu %zu %d %d\\n", sizeof(mpmc_b200_mc_move), sizeof(mpmc_b200_mc_result), MPMC_B200_MC_MAX_ATOMS, MPMC_B200_MC_MAX_STEPS);')
    for f in fields_m:
        body.append(f'printf("%zu\\n", offsetof(mpmc_b200_mc_move, {f}));')
    for f in fields_r:
        body.append(f'printf("%zu\\n", offsetof(mpmc_b200_mc_result, {f}));')
    body.append("return 0;}")
    src.write_text("\n".join(body))
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    out = subprocess.check_output([str(exe)], text=True).split()
    sizes = [int(x) for x in out[:3]]
    assert sizes == [ctypes.sizeof(api.Params), ctypes.sizeof(api.Energies), ctypes.sizeof(api.Counters)]
    n3 = len(fields_p) + len(fields_e) + len(fields_c)
    offs = [int(x) for x in out[3:3 + n3]]
    want = [getattr(api.Params, f).offset for f in fields_p] + [getattr(api.Energies, f).offset for f in fields_e] + [
        getattr(api.Counters, f).offset for f in fields_c]
    assert offs == want
    rest = [int(x) for x in out[3 + n3:]]
    assert rest[:4] == [ctypes.sizeof(api.McMove), ctypes.sizeof(api.McResult), api.MC_MAX_ATOMS, api.MC_MAX_STEPS]
    assert rest[4:] == [getattr(api.McMove, f).offset for f in fields_m] + [getattr(api.McResult, f).offset for f in fields_r]


def test_no_cpu_fallback(lib):
    """Without a CUDA device create() must fail with a message -- never fall back."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(api.EngineError) as ei:
        mpmc_b200.Engine(0)
    assert "no CUDA device" in str(ei.value) or "CUDA" in str(ei.value)


def test_product_does_not_import_oracle():
    """The oracle is test infrastructure: nothing under mpmc_b200/ may reference it."""
    bad = []
    for dp, _, files in os.walk(os.path.join(ROOT, "mpmc_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".c", ".cpp")) or f == "Makefile":
                text = open(os.path.join(dp, f), errors="ignore").read()
                if re.search(r"\boracle\b", text) and "test infrastructure" not in text.lower():
                    bad.append(os.path.join(dp, f))
    assert not bad, bad


def test_default_params_follow_reference_defaults():
    s = systems.h2_fluid(100)
    p = s.params
    assert p.ewald_kmax == 7 and p.polar_max_iter == 10 and p.polar_gamma == 1.0 and p.rd_lrc == 1
    L = p.basis[0][0]
    assert abs(p.cutoff - 0.5 * L) < 1e-12 and abs(p.ewald_alpha - 3.5 / p.cutoff) < 1e-15
    assert abs(p.reciprocal_basis[1][1] - 1.0 / L) < 1e-15 and abs(p.volume - L ** 3) < 1e-9 * L ** 3


def test_pqr_writer_roundtrip(tmp_path):
    """The PQR/input writers reproduce a System bit-for-bit through the reference parser
    (only where the reference build is available, i.e. in the build container)."""
    from oracle import oracle

    if not os.path.exists(oracle.REF_DRIVER):
        pytest.skip("oracle/_ref not built here")
    s = systems.mixed_small()
    inp = systems.write_reference_case(s, str(tmp_path))
    sr, exp, _ = oracle.run_reference(inp)
    assert np.array_equal(sr.pos, s.pos) and np.array_equal(sr.charge, s.charge)
    assert np.array_equal(sr.mol_id, s.mol_id) and np.array_equal(sr.frozen, s.frozen)
    assert sr.params.cutoff == s.params.cutoff and sr.params.ewald_alpha == s.params.ewald_alpha
