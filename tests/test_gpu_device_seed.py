"""The device-resident seed word (fpemu_set_seed_source / fpemu_advance_seed_source): stochastic kernels add it to the
host seed, so results equal the oracle at seed + word; a CUDA graph that contains the advance draws fresh random bits
on every replay; deterministic modes ignore it."""
import numpy as np
import pytest
import torch

from oracle import fpemu_oracle as O
from test_gpu_cast_parity import assert_same, dev, host, wide

pytestmark = pytest.mark.gpu
GOLDEN = 0x9E3779B97F4A7C15
MASK = (1 << 64) - 1


@pytest.fixture()
def F():
    import fpemu_b200
    fpemu_b200._lib.lib()
    yield fpemu_b200
    fpemu_b200.disable_device_seed()


def test_every_stochastic_entry_point_adds_the_device_word(F):
    O.set_nan_mode(False)
    x = wide(70003, 3)
    xt = dev(x)
    word = 123456789012345
    F.enable_device_seed(value=word)
    want = O.cast(x, "E4M3_STOCHASTIC", seed=(77 + word) & MASK)
    assert_same(host(F.cast(xt, "E4M3_STOCHASTIC", seed=77)), want, "cast")
    assert_same(host(F.cast(xt[1:], "E5M2_STOCHASTIC", seed=5, elem_offset=1)),
                O.cast(x[1:], "E5M2_STOCHASTIC", seed=(5 + word) & MASK, elem_offset=1), "generic (unaligned) cast")
    assert (F.encode(xt, "E3M4_STOCHASTIC", seed=9).cpu().numpy() == O.encode(x, "E3M4_STOCHASTIC", seed=(9 + word) & MASK)).all()
    n = 64 * 1024
    assert_same(host(F.cast_blocked(xt[:n], "E4M3_STOCHASTIC", 64, seed=11)),
                O.cast_blocked(x[:n], "E4M3_STOCHASTIC", 64, seed=(11 + word) & MASK), "blocked")
    x2 = x[:64 * 1024].reshape(64, 1024)
    got, _ = F.cast_rowwise(dev(x2), "E4M3_STOCHASTIC", seed=13, return_scales=True)
    s = [O.scale_from_amax(O.amax(r), 448.0) for r in x2]
    want2 = np.stack([O.cast(r, "E4M3_STOCHASTIC", scale=s[i], seed=(13 + word) & MASK, elem_offset=i * 1024) for i, r in enumerate(x2)])
    assert_same(host(got).reshape(-1), want2.reshape(-1), "row-wise")
    xs = [dev(wide(4096, 8)), dev(wide(33333, 9))]
    plan = F.MultiTensorPlan(xs, [torch.empty_like(t) for t in xs])
    outs = plan.cast("E5M2_STOCHASTIC", seed=100)
    for t, (xi, yi) in enumerate(zip(xs, outs)):
        assert_same(host(yi), O.cast(host(xi), "E5M2_STOCHASTIC", seed=(100 + t + word) & MASK), "multi %d" % t)
    y = F.cast_amax_scaled(xt, "E4M3_STOCHASTIC", seed=21)                       # single-launch path
    sc = O.scale_from_amax(O.amax(x), 448.0)
    assert_same(host(y), O.cast(x, "E4M3_STOCHASTIC", scale=sc, seed=(21 + word) & MASK), "fused amax-scaled")
    # deterministic modes do not look at the word
    assert_same(host(F.cast(xt, "E4M3_RNE")), O.cast(x, "E4M3_RNE"), "RNE")
    F.disable_device_seed()
    assert_same(host(F.cast(xt, "E4M3_STOCHASTIC", seed=77)), O.cast(x, "E4M3_STOCHASTIC", seed=77), "word removed")


def test_graph_replays_draw_fresh_bits(F):
    O.set_nan_mode(False)
    x = wide(1 << 16, 12)
    xt = dev(x)
    y = torch.empty_like(xt)
    F.enable_device_seed(value=1000)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        F.cast(xt, "E4M3_STOCHASTIC", out=y, seed=3)                               # warm-up (loads the kernels)
        F.advance_device_seed()
        s.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=s):
            F.cast(xt, "E4M3_STOCHASTIC", out=y, seed=3)
            F.advance_device_seed()
    word = (1000 + GOLDEN) & MASK                                                  # after the warm-up's advance
    outs = []
    for _ in range(3):
        graph.replay()
        torch.cuda.synchronize()
        assert_same(host(y), O.cast(x, "E4M3_STOCHASTIC", seed=(3 + word) & MASK), "replay")
        outs.append(y.clone())
        word = (word + GOLDEN) & MASK
    assert not torch.equal(outs[0].view(torch.int32), outs[1].view(torch.int32))
    assert not torch.equal(outs[1].view(torch.int32), outs[2].view(torch.int32))
    with pytest.raises(RuntimeError):
        F.disable_device_seed()
        F.advance_device_seed()
