"""fgdm_groupnorm_noalloc (no counterpart in the reference): the GroupNorm
whose scratch is a static __device__ array inside libfgdm_b200.so, for callers that cannot call cudaMalloc.

Called through the C-ABI (ctypes) and compared with the existing fgdm_groupnorm entry point on the same inputs: the same
two kernels run with the same geometry and fixed-order reductions, so the outputs are required to be BIT-equal where
fgdm_groupnorm takes its two-launch form, and within the GroupNorm tolerance of tests/test_kernels_gpu.py (2e-2 abs on
unit-variance outputs, bf16) where fgdm_groupnorm picks its one-launch slab kernel. Two calls in a row with different
shapes (the second reuses the scratch the first left behind). This file is named so that pytest collects it last.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

from fgdm_b200 import _lib  # noqa: E402

DEV = "cuda"
GN_TOL = 2e-2


def _p(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def _inputs(nb, hw, c0, c1, seed, mean=0.0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    x0 = (torch.randn(nb, hw, c0, generator=g) * 1.7 + mean).to(DEV).to(torch.bfloat16)
    x1 = (torch.randn(nb, hw, c1, generator=g) * 0.6 - mean).to(DEV).to(torch.bfloat16) if c1 else None
    gamma = (torch.randn(c0 + c1, generator=g) * 0.3 + 1.0).to(DEV)
    beta = (torch.randn(c0 + c1, generator=g) * 0.2).to(DEV)
    return x0, x1, gamma, beta


def _both(lib, nb, hw, c0, c1, groups, eps, silu, seed, mean=0.0):
    """(existing entry point with a caller workspace, new entry point without) on the same inputs."""
    x0, x1, gamma, beta = _inputs(nb, hw, c0, c1, seed, mean)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    ws = torch.empty(nb * 129 * groups * 2, device=DEV, dtype=torch.float32)
    y_ref = torch.full((nb, hw, c0 + c1), float("nan"), device=DEV, dtype=torch.bfloat16)
    y_new = torch.full((nb, hw, c0 + c1), float("nan"), device=DEV, dtype=torch.bfloat16)
    rc = lib.fgdm_groupnorm(_p(x0), _p(x1), _p(y_ref), _p(gamma), _p(beta), _p(ws), nb, hw, c0, c1, groups, eps, silu,
                            stream)
    assert rc == 0, lib.fgdm_last_error().decode()
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = lib.fgdm_groupnorm_noalloc(_p(x0), _p(x1), _p(y_new), _p(gamma), _p(beta), nb, hw, c0, c1, groups, eps, silu,
                                    stream)
    assert rc == 0, lib.fgdm_last_error().decode()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before          # nothing went through torch's allocator either
    # fp32 statement of GroupNorm(+SiLU) on the same bf16 inputs (util.py:223-225, openaimodel.py:221-224)
    x = torch.cat([x0] + ([x1] if x1 is not None else []), dim=-1).float()
    xn = torch.nn.functional.group_norm(x.permute(0, 2, 1), groups, gamma, beta, eps).permute(0, 2, 1)
    if silu:
        xn = torch.nn.functional.silu(xn)
    return y_ref, y_new, xn


# (nb, hw, c0, c1, groups, eps, silu, two_launch): two_launch marks shapes where fgdm_groupnorm itself runs the
# stats + apply pair (more than 16 slab rounds), so the comparison is bit-exact
CASES = [
    (16, 4096, 320, 0, 32, 1e-5, 1, True),       # UNet level 0 at the bench's R = 16 rows
    (4, 1024, 640, 320, 32, 1e-5, 1, True),      # skip concat, groups straddle the seam
    (2, 64, 1280, 0, 32, 1e-6, 0, False),        # 8x8 middle block: fgdm_groupnorm uses its one-launch slab form
    (64, 256, 128, 0, 32, 1e-6, 1, False),       # nb * groups = 2048: the stated maximum
]


def test_noalloc_matches_groupnorm_two_calls_in_a_row():
    _lib.require_device()
    lib = _lib.load()
    for rep in range(2):                          # the second pass runs on scratch the first pass filled
        for i, (nb, hw, c0, c1, groups, eps, silu, two_launch) in enumerate(CASES):
            y_ref, y_new, y32 = _both(lib, nb, hw, c0, c1, groups, eps, silu, seed=100 * rep + i, mean=3.0 * rep)
            assert not torch.isnan(y_new.float()).any()
            d = float((y_new.float() - y_ref.float()).abs().max())
            d32 = float((y_new.float() - y32).abs().max())
            print(f"noalloc GN pass {rep} nb={nb} hw={hw} c={c0}+{c1}: vs fgdm_groupnorm {d:.3e}, vs fp32 {d32:.3e}")
            if two_launch:
                assert torch.equal(y_new, y_ref)
            assert d <= GN_TOL
            assert d32 <= 2.5 * GN_TOL            # bf16 output rounding of values up to ~|6|


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    _lib.require_device()
    lib = _lib.load()
    x0, _, gamma, beta = _inputs(1, 16, 64, 0, 7)
    y = torch.empty_like(x0)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    # nb * groups = 65 * 32 > 2048; the kernel must not be launched (pointers cover one image only)
    rc = lib.fgdm_groupnorm_noalloc(_p(x0), None, _p(y), _p(gamma), _p(beta), 65, 16, 64, 0, 32, 1e-5, 0, stream)
    assert rc != 0
    msg = lib.fgdm_last_error().decode()
    assert "static scratch" in msg and "2048" in msg
    rc = lib.fgdm_groupnorm_noalloc(_p(x0), None, _p(y), _p(gamma), _p(beta), 257, 16, 64, 0, 4, 1e-5, 0, stream)
    assert rc != 0 and "256" in lib.fgdm_last_error().decode()
    torch.cuda.synchronize()
    # and a valid call right after an error still works
    y_ref, y_new, _ = _both(lib, 2, 4096, 320, 0, 32, 1e-5, 1, seed=9)
    assert torch.equal(y_new, y_ref)
