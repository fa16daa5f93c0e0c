"""bench.py's roofline numerator follows the algorithm that is executed: the Gram-delayed Lanczos form
on the multi-kernel workloads, the two-pass form in the persistent kernel and under NLS_REORTH=0; the reference line is
printed without a GPU."""
import importlib.util
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench():
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    return b


def test_bytes_per_site_step_follow_the_executed_form(monkeypatch):
    b = _bench()
    monkeypatch.delenv("NLS_REORTH", raising=False)
    W = b.WORKLOADS
    assert b.lz_impl(10) == 117 and b.lz_impl(12) == 165                  # two-pass: (m-1)(m+3)
    assert b.lz_impl(10, True) == 90 and b.lz_impl(12, True) == 121        # Gram-delayed: (m-1)(m+10)/2
    assert b.bytes_per_site_step(W["c5"]) == 16 * (90 + 10 + 2.5) == 1640
    assert b.bytes_per_site_step(W["c4"]) == 1640
    assert b.bytes_per_site_step(W["c3"]) == 16 * (121 + 12 + 2.5 + 1) == 2184
    assert b.bytes_per_site_step(W["c1"]) == 2072                          # persistent kernel: two-pass form
    assert b.bytes_per_site_step(W["c2_gautschi"]) == 2120
    assert b.bytes_per_site_step(W["c2_sv"]) == 32
    monkeypatch.setenv("NLS_REORTH", "0")                                  # A/B knob: the two-pass form everywhere
    assert b.bytes_per_site_step(W["c5"]) == 2072 and b.bytes_per_site_step(W["c3"]) == 2888
    # SURVEY's model is independent of the form
    assert b.bytes_per_site_step(W["c5"], model=True) == 2920


def test_every_workload_names_its_config():
    b = _bench()
    for name, w in b.WORKLOADS.items():
        assert {"desc", "n", "dim", "eq", "integ", "m", "dt", "seg"} <= set(w), name
