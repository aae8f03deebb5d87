"""Fixtures generated ON A B200 from the reference's own CUDA kernels (attention_kernels.cu, cache_kernels.cu,
pos_encoding_kernels.cu compiled unchanged for sm_100a) and its Triton prefix-prefill kernels
(tests/golden/make_golden_cuda.py -> tests/golden/ref_cuda_golden.npz).  They travel with the repo, so the pin holds on
every box whether or not oracle/_ref does:

  * CPU suite: the oracle (oracle/paged_attn_oracle.c) reproduces the reference's GPU outputs;
  * GPU suite: the product ops reproduce them too.
"""
import pytest
import torch

from tests import golden_cases
from tests.golden_io import load_cuda_golden

CASES = load_cuda_golden()
NAMES = sorted(CASES)


def test_fixture_covers_the_path():
    kinds = {n.split("_")[0] for n in NAMES}
    assert kinds == {"attn", "reshape", "convert", "rope", "prefill"}
    assert {"attn_bf16_fp8", "attn_f16_alibi_b32", "attn_f32_b8", "prefill_window_bf16", "prefill_alibi_f16"} <= set(NAMES)


@pytest.mark.parametrize("name", NAMES)
def test_oracle_reproduces_reference_cuda(oracle, name):
    golden_cases.check_case(oracle, name, CASES[name], "cpu")


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_product_reproduces_reference_cuda(ops, name):
    golden_cases.check_case(ops, name, CASES[name], "cuda:0")
