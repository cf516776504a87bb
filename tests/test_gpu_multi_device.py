"""Every entry point on a GPU that is NOT torch's current device (needs two GPUs; skipped otherwise): results must
equal those of device 0, the caller's current device must be left unchanged, and per-device caches (SM count,
resident-CTA count of the cooperative kernel, the no-alloc scratch address) must not leak across devices."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def F():
    import fpemu_b200
    fpemu_b200._lib.lib()
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    return fpemu_b200


def both(fn, *tensors):
    """fn on copies living on cuda:0 and on cuda:1, with cuda:0 current; returns the two results on the host."""
    torch.cuda.set_device(0)
    r0 = fn(*[t.to("cuda:0") for t in tensors])
    r1 = fn(*[t.to("cuda:1") for t in tensors])
    assert torch.cuda.current_device() == 0
    assert r1.device.index == 1
    torch.cuda.synchronize(0)
    torch.cuda.synchronize(1)
    return r0.cpu(), r1.cpu()


def same(a, b):
    return torch.equal(a.view(torch.uint8), b.view(torch.uint8))


def test_entry_points_on_a_non_current_device(F):
    x = torch.randn(300000 + 5) * 0.3
    x2 = torch.randn(96, 1024)
    for fn in (
        lambda t: F.cast(t, "E4M3_RNE"),
        lambda t: F.cast(t, "E5M2_STOCHASTIC", seed=9),
        lambda t: F.cast(t, "E4M3_RNE", d_scale=torch.tensor([3.7], device=t.device)),
        lambda t: F.cast_amax_scaled(t, "E4M3_RNE"),
        lambda t: F.cast_amax_scaled(t, "E4M3_RNE", fused=False),
        lambda t: F.cast_amax_scaled_noalloc(t, "E3M4_RNE", torch.empty_like(t)),
        lambda t: F.cast_blocked(t[:299008], "E4M3_RNE", 64),
        lambda t: F.encode(t, "E4M3_STOCHASTIC", seed=4),
        lambda t: F.decode(F.encode(t, "E5M2_RNE"), "E5M2"),
        lambda t: F.amax(t),
    ):
        a, b = both(fn, x)
        assert same(a, b)
    a, b = both(lambda t: F.cast_rowwise(t, "E4M3_RNE"), x2)
    assert same(a, b)
    a, b = both(lambda t: F.cast(t.to(torch.bfloat16), "E4M3_RNE"), x)
    assert same(a, b)


def test_multi_tensor_plan_on_a_non_current_device(F):
    torch.cuda.set_device(0)
    g = torch.Generator().manual_seed(3)
    host = [torch.randn(n, generator=g) for n in (4096, 70001, 33)]
    outs = []
    for dev in ("cuda:0", "cuda:1"):
        xs = [h.to(dev) for h in host]
        ys = [torch.empty_like(x) for x in xs]
        plan = F.MultiTensorPlan(xs, ys)
        plan.amax_scaled_cast("E4M3_RNE")
        torch.cuda.synchronize(torch.device(dev))
        outs.append([y.cpu() for y in ys])
    assert torch.cuda.current_device() == 0
    for a, b in zip(*outs):
        assert same(a, b)


def test_streams_of_two_devices_do_not_serialise_on_each_other(F):
    xs = [torch.randn(1 << 24, device="cuda:%d" % d) for d in (0, 1)]
    ys = [torch.empty_like(x) for x in xs]
    for _ in range(3):
        for x, y in zip(xs, ys):
            F.cast(x, "E4M3_RNE", out=y)
    for d in (0, 1):
        torch.cuda.synchronize(d)
    want = F.cast(xs[0], "E4M3_RNE").cpu()
    assert same(ys[0].cpu(), want)
    assert same(ys[1].cpu(), F.cast(xs[1], "E4M3_RNE").cpu())
