"""SQL -> planner -> plan IR checked on the CPU: the plans are executed by oracle/plan_interp.py (a plain-Python restatement of
the IR's operator semantics) and compared with the hand-written query restatements of oracle/nds_queries.py; and the SQL
test functions of the GPU suite (tests/test_gpu_sql_joins.py) are run with that interpreter standing in for the engine, so
their expectations are pinned by two independent executors - the engine on the GPU, the interpreter here."""
import math

import pyarrow as pa
import pytest

from oracle import nds_queries as oq
from oracle import plan_interp
from spark_rapids_benchmarks_b200 import planner, streams

PARAMS = [{'MANUFACT': 606, 'MANAGER': 63, 'MONTH': 1, 'YEAR': 1998, 'GEN': 'M', 'MS': 'S', 'ES': 'College'},
          {'MANUFACT': 17, 'MANAGER': 8, 'MONTH': 2, 'YEAR': 1998, 'GEN': 'F', 'MS': 'W', 'ES': 'Primary'}]


def _rows_equal(got, exp):
    if len(got) != len(exp):
        return False
    for g, e in zip(got, exp):
        if g.keys() != e.keys():
            return False
        for k in g:
            if isinstance(e[k], float) and g[k] is not None:
                if not math.isclose(g[k], e[k], rel_tol=1e-9):
                    return False
            elif g[k] != e[k]:
                return False
    return True


@pytest.mark.parametrize('query', streams.SUBSET + streams.EXTRA)
@pytest.mark.parametrize('pi', range(len(PARAMS)))
def test_planned_statements_match_the_oracle(warehouse, query, pi):
    sql = streams.render(query, overrides=PARAMS[pi])
    exp = oq.run_query(query, sql, warehouse)
    rb, _ = plan_interp.execute(planner.plan_sql(sql), warehouse)
    got = pa.Table.from_batches([rb])
    assert got.schema.names == exp.schema.names
    assert [str(t) for t in got.schema.types] == [str(t) for t in exp.schema.types], (got.schema, exp.schema)
    assert _rows_equal(got.to_pylist(), exp.to_pylist()), (got.to_pylist()[:3], exp.to_pylist()[:3])
    if pi == 0:
        assert exp.num_rows > 0


@pytest.mark.parametrize('query', streams.SUBSET)
def test_distributed_plan_forms_match_the_oracle(warehouse, query):
    """The one-plan forms of a distributed statement (partial aggregate -> exchange -> merge ...; gather and hash exchange) on one
    rank, where the exchange passes its child through: same rows as the single-GPU plan."""
    sql = streams.render(query, overrides=PARAMS[0])
    exp = oq.run_query(query, sql, warehouse).to_pylist()
    plan = planner.plan_sql(sql, distributed=True)
    for form in ('fused', 'fused_hash'):
        rb, _ = plan_interp.execute({'root': plan[form]}, warehouse)
        assert _rows_equal(rb.to_pylist(), exp), form


def test_sql_shapes_of_the_gpu_suite_on_the_interpreter(warehouse):
    from tests import test_gpu_sql_joins as m
    eng = plan_interp.FakeEngine(warehouse)
    names = [n for n in dir(m) if n.startswith('test_')]
    assert len(names) >= 12
    for n in names:
        getattr(m, n)(eng, warehouse)


def test_distributed_window_statement_keeps_the_gather_form(warehouse):
    """query98 across GPUs: partial aggregate -> exchange (gather) -> merge -> window -> ...; the hash form, where every rank
    would only see its own groups of a partition, is not offered."""
    sql = streams.render('query98')
    plan = planner.plan_sql(sql, distributed=True)
    assert 'fused' in plan and 'fused_hash' not in plan
    rb, _ = plan_interp.execute({'root': plan['fused']}, warehouse)
    assert _rows_equal(rb.to_pylist(), oq.run_query('query98', sql, warehouse).to_pylist())


@pytest.mark.parametrize('world', [2, 3])
@pytest.mark.parametrize('query', streams.SUBSET + streams.EXTRA)
def test_distributed_plan_forms_over_emulated_ranks(warehouse, query, world):
    """Every rank of `world` runs the one-plan form over its shard of the fact files (every world-th file), the exchange gathers
    (or hash-partitions) the partial states of all ranks: every rank must end with the single-GPU answer."""
    sql = streams.render(query, overrides=PARAMS[0])
    exp = oq.run_query(query, sql, warehouse).to_pylist()
    plan = planner.plan_sql(sql, distributed=True)
    for form in ('fused', 'fused_hash'):
        if form not in plan:
            continue
        for rank in range(world):
            rb, _ = plan_interp.execute({'root': plan[form]}, warehouse, rank, world)
            assert _rows_equal(rb.to_pylist(), exp), (form, rank)
    # a statement whose root table is replicated is planned without an exchange: every rank computes it alone
    rep = planner.plan_sql('select i_category, count(*) n from item group by i_category order by i_category', distributed=True)
    assert rep.get('replicated') and 'fused' not in rep
