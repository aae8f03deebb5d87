import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box: pytest -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    """CPU oracle (oracle/paged_attn_oracle.c) — the checker, never the thing under test."""
    from oracle import oracle as _oracle
    _oracle.build()
    return _oracle


@pytest.fixture(scope="session")
def golden():
    from tests.golden_io import load_golden
    return load_golden()


@pytest.fixture(scope="session")
def ref_cpu():
    from tests import helpers
    mod = helpers.load_ref_cpu()
    if mod is None:
        pytest.skip("oracle/_ref/ref_cpu*.so not built or host lacks AVX512")
    return mod


@pytest.fixture(scope="session")
def ref_cuda():
    from tests import helpers
    mod = helpers.load_ref_cuda()
    if mod is None:
        pytest.skip("oracle/_ref/ref_cuda.so not built")
    return mod


@pytest.fixture(scope="session")
def ref_cuda_optional():
    """ref_cuda when built, else None (for tests whose main checker is the CPU oracle)."""
    from tests import helpers
    return helpers.load_ref_cuda()


@pytest.fixture(scope="session")
def ops():
    """The product ops (CUDA extension).  Import fails loudly when the extension is missing."""
    from vllm_legacy_b200 import _custom_ops
    return _custom_ops
