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


# The tests exercise THIS repository's kernels behind torch.ops._C / torch.ops._C_cache_ops.  Upstream vLLM's compiled op
# libraries define the same op names; a process can hold both only when upstream's is loaded FIRST (the plugin order:
# our shim then overrides the CUDA kernels of the identical schemas, csrc/torch_shim.cpp register_op) — loaded second,
# upstream's TORCH_LIBRARY definition aborts the process.  tests/test_v1_backend.py imports upstream's pure-Python
# interface modules, and on a CUDA box `vllm.platforms.cuda` does `import vllm._C` as a side effect, so the compiled
# modules are replaced by empty stand-ins for the test session: no upstream kernel can run in these tests.
# (tools/upstream_witness.py, the one place upstream's kernels are wanted, runs in its own process.)
import types  # noqa: E402

for _name in ("vllm._C", "vllm._C_stable_libtorch"):
    sys.modules.setdefault(_name, types.ModuleType(_name))


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
