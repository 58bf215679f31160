"""Files from outside (results CSVs, MatrixMarket matrices) that are damaged must produce an
error, not a crash or a silently wrong object."""
import pytest
from hypothesis import given, settings, strategies as st

import tenzing_b200 as tz
from tenzing_b200 import _C
from tenzing_b200.models import DistributedSpMV


def _graph():
    g = tz.Graph()
    for n in ("a", "b"):
        o = tz.NoOp(n)
        g.start_then(o)
        g.then_finish(o)
    return g


def test_damaged_results_csv():
    g = _graph()
    plat = tz.Platform.make_null(1)
    res = tz.dfs.explore(g, plat, tz.PyBenchmarker(lambda s: 1.0))
    good = res.csv()
    assert tz.CsvBenchmarker.from_text(good, g).num_rows() == 2
    lines = good.splitlines()
    # an op the graph does not know
    with pytest.raises(RuntimeError):
        tz.CsvBenchmarker.from_text(lines[1].replace('{"kind":"NoOp","name":"a"}', '{"name":"zzz"}'), g)
    # broken JSON in an op column
    with pytest.raises(RuntimeError):
        tz.CsvBenchmarker.from_text(lines[1].replace('{"kind":"NoOp","name":"a"}', '{"name":'), g)
    # a statistics column that is not a number, after valid rows
    with pytest.raises(RuntimeError):
        tz.CsvBenchmarker.from_text(good + lines[1].replace("|", "|x", 1).replace("0|", "7|", 1) + "\n", g)
    # short / empty / comment lines are skipped
    assert tz.CsvBenchmarker.from_text("# comment\n\n1|2|3\n" + good, g).num_rows() == 2


@settings(max_examples=150, deadline=None)
@given(st.text(alphabet='0123456789|{}":,.e-namStrtFiish ab\n', max_size=120))
def test_random_csv_text_never_crashes(text):
    try:
        tz.CsvBenchmarker.from_text(text, _graph())
    except RuntimeError:
        pass


@pytest.mark.parametrize("content", [
    "",                                                             # empty
    "%%MatrixMarket matrix array real general\n2 2\n1\n2\n3\n4\n",  # dense format
    "%%MatrixMarket matrix coordinate real general\n3 3 2\n1 1 1.0\n",        # fewer entries than promised
    "%%MatrixMarket matrix coordinate real general\n3 3 1\n7 1 1.0\n",        # row out of range
    "%%MatrixMarket matrix coordinate real general\n3 4 1\n1 1 1.0\n",        # not square
    "%%MatrixMarket matrix coordinate real general\nx y z\n",                 # size line is not numbers
    "hello\n1 1 1\n1 1 1.0\n",                                               # no banner
])
def test_damaged_matrix_market_files_are_refused(tmp_path, content):
    path = tmp_path / "m.mtx"
    path.write_text(content)
    with pytest.raises((RuntimeError, ValueError)):
        DistributedSpMV(matrix_file=str(path), host=True)


def test_good_matrix_market_variants_are_read(tmp_path):
    path = tmp_path / "m.mtx"
    path.write_text("%%MatrixMarket matrix coordinate real symmetric\n% comment\n3 3 3\n1 1 2.0\n2 1 1.0\n3 3 4.0\n")
    s = DistributedSpMV(matrix_file=str(path), host=True)
    assert s.spmv.global_rows() == 3 and s.spmv.local_nnz() + s.spmv.remote_nnz() == 4   # mirrored entry
    path.write_text("%%MatrixMarket matrix coordinate pattern general\n2 2 2\n1 2\n2 1\n")
    s = DistributedSpMV(matrix_file=str(path), host=True)
    assert s.spmv.local_nnz() == 2


@pytest.mark.timeout(60)
def test_a_bad_matrix_file_fails_on_every_rank_instead_of_hanging_the_others(tmp_path):
    path = tmp_path / "bad.mtx"
    path.write_text("%%MatrixMarket matrix coordinate real general\n3 3 1\n9 9 1.0\n")
    seen = {}

    def body(rank, comm):
        try:
            DistributedSpMV(comm=comm, matrix_file=str(path), host=True)
        except RuntimeError as e:
            seen[rank] = str(e)

    _C.loopback_run(3, body)
    assert set(seen) == {0, 1, 2} and all("outside the 3 x 3 matrix" in m for m in seen.values())


def test_workload_arguments_are_validated():
    from tenzing_b200.models import HaloExchange
    for kw in (dict(nX=2, nGhost=3), dict(nX=0), dict(nX=4, nQ=0), dict(nX=4, nGhost=-1),
               dict(nX=4, nGhost=1, neighbors=7)):
        with pytest.raises(RuntimeError):
            HaloExchange(host=True, **{"nQ": 1, "nGhost": 1, **kw})
    with pytest.raises(ValueError):
        HaloExchange(nX=4, nQ=1, nGhost=1, host=True, rank_grid=(2, 1, 1))     # one rank here
    # the ghost layer may be as wide as the interior, and it may be absent
    assert HaloExchange(nX=3, nQ=1, nGhost=3, host=True).domain.extent() == [9, 9, 9]
    assert HaloExchange(nX=3, nQ=1, nGhost=0, host=True).bytes_per_exchange() == 0


@pytest.mark.timeout(120)
def test_more_ranks_than_matrix_rows():
    """ranks that own no rows still take part in the handshake of every style"""
    from tenzing_b200.models import common
    from tenzing_b200.models import spmv as spmv_mod

    def body(rank, comm):
        for n in (2, 5):
            s = DistributedSpMV(comm=comm, n=n, nnz_per_row=3, host=True)
            s.platform.set_host_execution(True)
            for style in spmv_mod.STYLES + spmv_mod.EXPERIMENTAL_STYLES:
                seq = common.fixed_schedule(s.graph(style), s.platform)
                s.fill()
                comm.barrier()
                s.run(seq, 2)
                comm.barrier()
                assert s.check(), (n, style, rank)
                comm.barrier()

    _C.loopback_run(4, body)


def test_serialized_ops_with_absurd_ids_are_refused():
    g = _graph()
    for op in ('{"event":2000000000,"kind":"CudaEventRecord","name":"x","stream":0}',
               '{"event":-3,"kind":"CudaEventSync","name":"x"}',
               '{"event":0,"kind":"CudaStreamWaitEvent","name":"x","stream":-1}'):
        with pytest.raises(RuntimeError):
            _C.sequence_from_json('[{"name":"Start"},' + op + ',{"name":"Finish"}]', g)
