"""llm-inference_b200 — B200-native (sm_100a) LLaMA decoder hot path.

The product is the C-ABI shared library ``lib/libb200llama.so`` (``include/b200_llama.h``) plus
the C++ header shim / layer classes under ``include/`` and ``host/``. This Python package is
plumbing only: it loads the library with ctypes and passes ``torch`` device pointers through it
(tests, bench). There is no CPU fallback: importing works anywhere, but every op raises if the
library is missing or no CUDA device is present.

The directory name carries a hyphen (it mirrors the reference's repository name), so import it
with ``__graft_entry__.load_package()`` / ``tests/conftest.py`` which register it as
``llm_inference_b200``.
"""
from . import capi  # noqa: F401
from .capi import lib, lib_path, F32, F16  # noqa: F401
from . import ops  # noqa: F401
from . import engine  # noqa: F401
from . import sharding  # noqa: F401

__all__ = ["capi", "ops", "engine", "sharding", "lib", "lib_path", "F32", "F16"]
