"""B200-native (sm_100a) paged-attention hot path of DeepAuto-AI/vllm-legacy (vLLM 0.4.3 fork).

Layout mirrors the reference for this path only:

  vllm_legacy_b200._C                      <->  vllm._C            (ops, cache_ops, cuda_utils)
  vllm_legacy_b200._custom_ops             <->  vllm/_custom_ops.py
  vllm_legacy_b200.attention.ops.paged_attn<->  vllm/attention/ops/paged_attn.py
  vllm_legacy_b200.attention.backends.*    <->  vllm/attention/backends/abstract.py + a decode backend

The compiled extension (`_C.so`, over the C-ABI library `libb200pa.so`, see include/b200pa.h) is
required: there is no CPU or eager fallback.  Build it with `python -m vllm_legacy_b200.build` or
`__graft_entry__.build()`.
"""
import os

__version__ = "0.1.0"

PACKAGE_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PACKAGE_DIR, "libb200pa.so")


def load_extension():
    """Import and return the compiled `_C` module; raise loudly if it is missing."""
    try:
        from . import _C  # noqa: F401
    except ImportError as e:  # pragma: no cover - exercised only on a broken checkout
        raise ImportError(
            "vllm_legacy_b200._C (the sm_100a CUDA extension) is not built or failed to load: "
            f"{e}. Run `python -m vllm_legacy_b200.build`. There is no fallback path.") from e
    return _C
