"""fpemu_cast_amax_scaled_noalloc — the entry point for callers that cannot allocate device memory (a plugin
in a server whose pool belongs to the host framework, code that owns no allocator, loops where cudaMalloc is
forbidden because it synchronises the device and can fail).  Its scratch is a static device array inside the
library, shared by every call on the GPU, so the test makes TWO calls in a row with different data, sizes
and element types and checks each against the regular caller-scratch path
(fpemu_amax + fpemu_scale_from_amax + fpemu_cast) and against the CPU oracle — bit for bit, NaN == NaN, the
tolerance every deterministic test here uses — and checks the property in the entry point's name: the free
device memory the driver reports does not change across the calls.  This file sorts last on purpose: a run
that stops at the first failure has then already executed every other test."""
import pytest
import torch

from oracle import fpemu_oracle as O

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


def _oracle(x, mode, flt_max):
    O.set_nan_mode(False)
    xn = x.cpu().numpy()
    return O.cast(xn, mode, scale=O.scale_from_amax(O.amax(xn), flt_max, like=xn))


def test_two_calls_in_a_row_match_the_caller_scratch_path(F):
    L = F._lib
    torch.manual_seed(123)
    # call 1: large fp32 tensor with a ragged tail, big dynamic range (fills many scratch slots)
    x1 = torch.randn((1 << 22) + 13, device="cuda") * 517.0
    # call 2: a much smaller fp16 tensor with a tiny maximum — stale slots of call 1 must not leak into its scale
    x2 = (torch.randn(50001, device="cuda") * 1e-3).half()
    outs = [torch.empty_like(x1), torch.empty_like(x2)]
    want = [F.cast_amax_scaled(x, mode, return_scale=True) for x, mode in ((x1, "E4M3_RNE"), (x2, "E5M2_RNE"))]
    stream = torch.cuda.current_stream().cuda_stream
    # a first call on this device (the one-time address look-up) may already have happened in this process or not
    # — make one, then measure: straight through the C ABI, outputs preallocated, nothing of torch's in between
    assert L.lib().fpemu_cast_amax_scaled_noalloc(x2.data_ptr(), outs[1].data_ptr(), 16, L.DT_F16, L.FMT_E5M2,
                                                  L.RND_RNE, 0, 57344.0, 0, 0, stream) == 0
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    for (x, y, dt, fmt, flt_max) in ((x1, outs[0], L.DT_F32, L.FMT_E4M3, 448.0), (x2, outs[1], L.DT_F16, L.FMT_E5M2, 57344.0)):
        st = L.lib().fpemu_cast_amax_scaled_noalloc(x.data_ptr(), y.data_ptr(), x.numel(), dt, fmt, L.RND_RNE, 0,
                                                    flt_max, 0, 0, stream)
        assert st == 0, L.strerror(st)
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before, "the no-alloc entry point changed the device's free memory"
    for (x, mode, flt_max), y, (w, w_scale) in zip(((x1, "E4M3_RNE", 448.0), (x2, "E5M2_RNE", 57344.0)), outs, want):
        assert _same(y, w), mode                                                     # vs the caller-scratch path
        assert torch.isfinite(w_scale).all()
        assert O.same_bits_or_both_nan(y.cpu().numpy(), _oracle(x, mode, flt_max)).all(), mode + " vs oracle"
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


def test_every_gpu_has_its_own_scratch(F):
    """The static array exists once per GPU that loads the library; its address is looked up per device.  With two
    GPUs: interleaved calls on both give what each gives alone."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    xs = [torch.randn(100003, device="cuda:%d" % d) * (3.0 + 100.0 * d) for d in (0, 1)]
    want = [F.cast_amax_scaled(x, "E4M3_RNE") for x in xs]
    got = [F.cast_amax_scaled_noalloc(x, "E4M3_RNE", out=torch.empty_like(x)) for x in xs]
    for d in (0, 1):
        torch.cuda.synchronize(d)
        assert _same(got[d], want[d])
