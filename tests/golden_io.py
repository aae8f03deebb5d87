"""Loader for tests/golden/ref_cpu_golden.npz (made by tests/golden/make_golden.py from the
reference's own CPU code).  Returns {case: {field: torch tensor}} with bf16 restored."""
import os
from collections import defaultdict

import numpy as np
import torch

PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_cpu_golden.npz")


def load_golden():
    z = np.load(PATH)
    cases = defaultdict(dict)
    for key in z.files:
        case, field = key.split(".", 1)
        cases[case][field] = z[key]
    out = {}
    for case, fields in cases.items():
        meta = fields["meta"]
        is_bf16 = bool(meta[-2]) if case.startswith("attn") else bool(meta[-1])
        d = {}
        for f, a in fields.items():
            t = torch.from_numpy(np.ascontiguousarray(a))
            if a.dtype == np.int16 and is_bf16:
                t = t.view(torch.bfloat16)
            d[f] = t
        out[case] = d
    return out


CUDA_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_cuda_golden.npz")
_DT = {"f16": torch.float16, "bf16": torch.bfloat16, "f32": torch.float32}


def load_cuda_golden():
    """tests/golden/ref_cuda_golden.npz (made ON A B200 by tests/golden/make_golden_cuda.py from the reference's own
    CUDA kernels and Triton prefix-prefill kernels).  {case: {field: tensor | str}}; 16-bit floats restored from their
    int16 bit patterns according to the case's `dtype`."""
    z = np.load(CUDA_PATH)
    cases = defaultdict(dict)
    for key in z.files:
        case, field = key.split(".", 1)
        cases[case][field] = z[key]
    out = {}
    for case, fields in cases.items():
        dt = _DT[str(fields["dtype"])]
        d = {}
        for f, a in fields.items():
            if a.dtype.kind in "US":
                d[f] = str(a)
                continue
            t = torch.from_numpy(np.ascontiguousarray(a))
            if a.dtype == np.int16:
                t = t.view(dt)
            d[f] = t
        d["torch_dtype"] = dt
        out[case] = d
    return out
