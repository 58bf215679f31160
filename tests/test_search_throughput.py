"""The search-throughput comparison (scripts/search_throughput.py): this
framework's enumeration counts are the closed form, and - when the reference has been built
(baseline/build_ref.py) - the reference's own `dfs::get_all_sequences` returns the same counts
on the same graphs, so the two timings in profiles/search_throughput.txt are of the same work."""
import importlib.util
import math
import os

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _script():
    spec = importlib.util.spec_from_file_location("search_throughput", os.path.join(ROOT, "scripts", "search_throughput.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.parametrize("chains,length", [(3, 1), (3, 2), (2, 3), (4, 2)])
def test_enumeration_counts_are_the_closed_form(chains, length):
    n, seconds = _script().ours(chains, length, repeats=1)
    assert n == math.factorial(chains * length) // math.factorial(length) ** chains
    assert seconds > 0


def test_reference_enumerates_the_same_number_of_schedules():
    st = _script()
    if not os.path.exists(os.path.join(ROOT, "baseline", "_ref", "build_e2e", "ref-enum")):
        pytest.skip("reference not built (python baseline/build_ref.py)")
    for chains, length in [(3, 1), (3, 2)]:
        n_ref, t_ref = st.reference(chains, length, repeats=1, limit_s=120)
        if n_ref is None:
            pytest.skip("the reference binary does not run on this machine (built for another CPU)")
        assert n_ref == st.ours(chains, length, repeats=1)[0]
        assert t_ref > 0
