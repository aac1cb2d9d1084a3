import os
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def emu_parser():
    """trace_parser over the SIMT-emulated build of the kernel source (CPU, test-only)."""
    import harness
    from accel_sim_framework_distributed_b200 import trace_parser
    tp = trace_parser(lib_path=harness.emu_library())
    yield tp
    tp.close()


@pytest.fixture(scope="session")
def cuda_parser():
    """trace_parser over the product CUDA library.  Fails (does not skip) without a GPU."""
    import harness
    from accel_sim_framework_distributed_b200 import trace_parser
    tp = trace_parser(lib_path=harness.CUDA_LIB)
    yield tp
    tp.close()
