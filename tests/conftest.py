import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_sessionfinish(session, exitstatus):
    """GPU sessions leave the actual counts of every kernel-vs-oracle comparison in profiles/parity_r02.json
    (and gpurun_out/, which is what travels back from the GPU box)."""
    try:
        sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
        import parity
        parity.write_report()
    except Exception as e:      # the report must never turn a green run red
        print("parity report not written:", e)


@pytest.fixture(scope="session")
def port_oracle():
    import oracle
    oracle.build_port()
    return oracle


@pytest.fixture(scope="session")
def rt_ctx():
    """One rt_context on cuda:0 through the C ABI. Fails (does not skip) when the CUDA library
    or the device is missing: GPU tests must never pass on a fallback."""
    from opengl_raytracing_b200.pipeline import RtContext
    ctx = RtContext(0)
    yield ctx
    ctx.close()
