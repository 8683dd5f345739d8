import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box with -m gpu)')


@pytest.fixture(scope='session')
def warehouse(tmp_path_factory):
    """Tiny seeded warehouse (SF 0.01, 40 sales days) shared by the whole session."""
    from spark_rapids_benchmarks_b200 import datagen
    d = os.environ.get('NDS_TEST_WAREHOUSE') or str(tmp_path_factory.mktemp('nds_sf001'))
    datagen.ensure(d, 0.01, sales_days=40, workers=1)
    return d


@pytest.fixture(scope='session')
def gpu_engine(warehouse):
    """Engine with every table of the tiny warehouse registered (GPU tests only)."""
    from spark_rapids_benchmarks_b200 import engine, nds_schema
    eng = engine.Engine(device=0, resident=True, keep_host=True)
    for t in nds_schema.get_schemas(True):
        eng.register_table(t, os.path.join(warehouse, t))
    yield eng
    eng.close()
