import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SOLVERS = os.path.join(ROOT, "robot_mpcs_b200", "solvers")
CONFIG = os.path.join(ROOT, "robot_mpcs_b200", "config")
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built():
    """Solver directories (CUDA .so + host twin); built on demand, normally prebuilt by __graft_entry__.build()."""
    import __graft_entry__ as ge
    missing = []
    for cfg in ge.CONFIGS:
        from robot_mpcs_b200.utils.utils import parse_setup
        s = parse_setup(os.path.join(CONFIG, cfg))["mpc"]
        name = f"{s['model_name']}_n{s['n']}_{str(s['time_step']).replace('.', '')}_H{s['time_horizon']}_noSlack"
        if not os.path.isfile(os.path.join(SOLVERS, name, f"lib{name}.so")) or \
                not os.path.isfile(os.path.join(SOLVERS, name, f"libtwin_{name}.so")):
            missing.append(cfg)
    if missing:
        ge.build()
    return SOLVERS
