"""B200-native decode hot path of vLLM-metax: paged_attention_v1/v2 and the KV-cache writers.

Importing the package does not load native code.  `import_kernels()` loads the two shared
objects built in-tree by `vllm_metax_b200.build`:

  libb200_paged_attn.so  the C ABI (include/b200_paged_attn.h) holding the sm_100a kernels
  _C.so                  the torch shim that registers torch.ops._C.paged_attention_v1/v2 and
                         torch.ops._C_cache_ops.reshape_and_cache(_flash)/convert_fp8

This mirrors the reference, where `MacaPlatform.import_kernels()` imports `vllm_metax._C`
to populate the same namespaces (reference: vllm_metax/platform.py:107-115,
csrc/torch_bindings.cpp:19,376).  There is no CPU or PyTorch fallback: if the libraries are
missing, `import_kernels()` raises.
"""
from __future__ import annotations

import importlib
import threading

__version__ = "0.1.0"

_lock = threading.Lock()
_loaded = False


def import_kernels() -> None:
    """Load libb200_paged_attn.so and register the torch ops. Raises if the build is missing."""
    global _loaded
    with _lock:
        if _loaded:
            return
        from . import _native

        _native.load()  # raises with a build hint if the .so is absent
        try:
            importlib.import_module("vllm_metax_b200._C")
        except ImportError as e:  # no silent fallback
            raise ImportError(
                "vllm_metax_b200._C (torch shim) is not built or failed to load: %s. "
                "Run `python -m vllm_metax_b200.build`." % e
            ) from e
        _loaded = True


def kernels_loaded() -> bool:
    return _loaded
