"""fpemu_cast_amax_scaled_noalloc — the entry point for callers that cannot allocate device memory (a plugin
in a server whose pool belongs to the host framework, code that owns no allocator, loops where cudaMalloc is
forbidden because it synchronises the device and can fail).  Its scratch is a static device array inside the
library, shared by every call on the GPU, so the test makes TWO calls in a row with different data, sizes
and element types and checks each against the regular caller-scratch path
(fpemu_amax + fpemu_scale_from_amax + fpemu_cast) — bit for bit, NaN == NaN, the tolerance every
deterministic test here uses.  This file sorts last on purpose: a run that stops at the first failure has
then already executed every other test."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def F():
    import fpemu_b200
    fpemu_b200._lib.lib()
    return fpemu_b200


def _bits(t):
    if t.dtype == torch.float32:
        return t.view(torch.int32)
    return t.view(torch.int16)


def _same(a, b):
    return bool(((_bits(a) == _bits(b)) | (a.isnan() & b.isnan())).all().item())


def test_two_calls_in_a_row_match_the_caller_scratch_path(F):
    torch.manual_seed(123)
    # call 1: large fp32 tensor with a ragged tail, big dynamic range (fills many scratch slots)
    x1 = torch.randn((1 << 22) + 13, device="cuda") * 517.0
    # call 2: a much smaller fp16 tensor with a tiny maximum — stale slots of call 1 must not leak into its scale
    x2 = (torch.randn(50001, device="cuda") * 1e-3).half()
    for x, mode in ((x1, "E4M3_RNE"), (x2, "E5M2_RNE")):
        want, want_scale = F.cast_amax_scaled(x, mode, return_scale=True)
        got = F.cast_amax_scaled_noalloc(x, mode, out=torch.empty_like(x))
        assert _same(got, want), mode
        assert torch.isfinite(want_scale).all()
    # and once more in the other order, in place, on an unaligned view and a bf16 tensor
    y = x2.clone()
    F.cast_amax_scaled_noalloc(y, "E3M4_RNE", out=y)
    assert _same(y, F.cast_amax_scaled(x2, "E3M4_RNE"))
    v = x1[3:100003]
    assert _same(F.cast_amax_scaled_noalloc(v, "E4M3_RNE", out=torch.empty_like(x1)[3:100003]), F.cast_amax_scaled(v, "E4M3_RNE"))
    xb = x1[: 1 << 20].bfloat16()
    assert _same(F.cast_amax_scaled_noalloc(xb, "E4M3_STOCHASTIC", out=torch.empty_like(xb), seed=5),
                 F.cast_amax_scaled(xb, "E4M3_STOCHASTIC", seed=5))


def test_sizes_beyond_the_static_capacity_are_rejected_without_a_launch(F):
    L = F._lib
    x = torch.ones(64, device="cuda")
    too_many = (1 << 27) + 1                                   # fp32 limit is 2^27 elements (512 MiB)
    st = L.lib().fpemu_cast_amax_scaled_noalloc(x.data_ptr(), x.data_ptr(), too_many, L.DT_F32, L.FMT_E4M3, L.RND_RNE, 0,
                                                448.0, 0, 0, None)
    assert st == L.ERR_TOO_LARGE
    assert "capacity" in L.strerror(st)
    st = L.lib().fpemu_cast_amax_scaled_noalloc(x.data_ptr(), x.data_ptr(), 64, L.DT_F32, L.FMT_BF16, L.RND_RNE, 0,
                                                448.0, 0, 0, None)
    assert st == L.ERR_MODE
    assert np.float32(torch.ones(1).item()) == 1.0
