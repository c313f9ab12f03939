"""pytest configuration: the `gpu` marker and the repo root on sys.path.

`-m "not gpu"` covers the oracle against the golden vectors, the host logic and that the C-ABI
library loads and exports what include/b200_paged_attn.h declares.  `-m gpu` tests are the parity
tests proper: CUDA path vs oracle on the same seeded inputs.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def ops():
    """The op wrappers; importing them loads the native libraries (raises if they are not built)."""
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from vllm_metax_b200 import _custom_ops

    return _custom_ops
