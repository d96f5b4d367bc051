"""Host logic of bench.py that needs no GPU: workload names with option suffixes, and that the CPU arm's bounded sample
carries the same options as the workload it stands for."""
import importlib.util
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_spec = importlib.util.spec_from_file_location("bench_module", os.path.join(ROOT, "bench.py"))
bench = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(bench)


def test_split_workload_peels_option_suffixes():
    assert bench.split_workload("mof_h2_63k") == ("mof_h2_63k", [])
    assert bench.split_workload("mof_h2_63k_ewald_full") == ("mof_h2_63k", ["_ewald_full"])
    assert bench.split_workload("co2_fluid_12k_es_wolf_rd_crystal2") == ("co2_fluid_12k", ["_es_wolf", "_rd_crystal2"])
    assert bench.split_workload("h2_fluid_4000") == ("h2_fluid_4000", [])


def test_workload_options_reach_the_system_and_the_cpu_sample():
    s = bench.build_workload("co2_fluid_12k_rd_crystal3")
    assert s.n == 12288 and s.params.rd_crystal == 1 and s.params.rd_crystal_order == 3
    s = bench.build_workload("co2_fluid_12k_es_wolf")
    assert s.params.es_wolf == 1 and s.params.feynman_hibbs == 0
    s = bench.build_workload("mof_h2_7k_gs_production")
    assert s.params.polar_gs_ranked == 1 and s.params.polar_max_iter == 4 and s.params.polar_palmo == 1
    plain, _ = bench._cpu_sample_system("co2_fluid_12k", 25.0, 3)
    sample, what = bench._cpu_sample_system("co2_fluid_12k_rd_crystal2", 25.0, 3)
    assert sample.params.rd_crystal == 1 and sample.params.rd_crystal_order == 2 and "_rd_crystal2" in what
    assert sample.n <= plain.n  # the image sums cost more per call: the sample shrinks, it never grows
    sample, _ = bench._cpu_sample_system("mof_h2_63k_ewald_full", 25.0, 3)
    assert sample.params.polar_ewald_full == 1
    sample, _ = bench._cpu_sample_system("mof_h2_63k", 25.0, 3)
    assert sample.params.polar_ewald_full == 0 and sample.params.rd_crystal == 0
