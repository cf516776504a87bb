"""The host-buffer entry points (fpemu_cast_host, fpemu_cast_host_multi): pinned host tensor in, pinned host tensor(s)
out, staged through device buffers in chunks.  Results must equal the device-resident casts (and the oracle) bit for
bit whatever the chunking — stochastic rounding included, because every chunk carries its global element offset."""
import numpy as np
import pytest
import torch

from oracle import fpemu_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def F():
    import fpemu_b200
    fpemu_b200._lib.lib()
    return fpemu_b200


def bits(t):
    return t.view(torch.int32 if t.dtype == torch.float32 else torch.int16)


def same(a, b):
    return bool(((bits(a) == bits(b)) | (a.isnan() & b.isnan())).all().item())


def wide(n, seed, dtype):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, generator=g) * torch.exp2(torch.randint(-12, 10, (n,), generator=g).float())
    x[:4] = torch.tensor([float("inf"), -0.0, 65520.0, 1e-40])
    return x.to(dtype).pin_memory()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
def test_host_cast_equals_the_device_cast_for_any_chunking(F, dtype):
    n = 300000 + 7                                              # ragged: the last chunk is short and ends off a vector
    h_in = wide(n, 1, dtype)
    x = h_in.cuda()
    for chunk in (1 << 12, 1 << 16, 1 << 20):                   # 74 chunks, 5 chunks, one chunk
        hc = F.HostCaster(dtype, chunk_elems=chunk)
        for mode, scale, seed in (("E4M3_RNE", 1.0, 0), ("E5M2_STOCHASTIC", 1.0, 11), ("E3M4_RNE", 3.7, 0),
                                  ("BFLOAT16_STOCHASTIC", 1.0, 5), ("E4M3_IEEE_RNE", 1.0, 0)):
            h_out = torch.empty(n, dtype=dtype).pin_memory()
            hc.run(h_in, h_out, mode, scale=scale, seed=seed)
            assert same(h_out, F.cast(x, mode, scale=scale, seed=seed).cpu()), (mode, chunk)
    O.set_nan_mode(False)
    if dtype == torch.float32:
        h_out = torch.empty(n).pin_memory()
        F.HostCaster(dtype, chunk_elems=1 << 14).run(h_in, h_out, "E4M3_RNE", scale=3.7)
        assert O.same_bits_or_both_nan(h_out.numpy(), O.cast(h_in.numpy(), "E4M3_RNE", scale=3.7)).all()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_one_upload_many_casts(F, dtype):
    n = 200000 + 24
    h_in = wide(n, 2, dtype)
    h_in[:4] = torch.tensor([3.0, -0.0, 100.0, 1e-3]).to(dtype)                 # finite: the amax-scaled jobs need a maximum
    x = h_in.cuda()
    modes = [("E5M2_RNE", 1.0, 0), ("E4M3_STOCHASTIC", 1.0, 9), ("E3M4_RTZ", 0.125, 0), ("BFLOAT16_RNE", 1.0, 0),
             ("E4M3_RNE", "amax", 0), ("FLOAT16_STOCHASTIC", 1.0, 3), ("E5M2_STOCHASTIC", "amax", 21), ("FP4_NEAREST", 1.0, 0)]
    for chunk in (1 << 13, 1 << 18):
        hc = F.HostCaster(dtype, chunk_elems=chunk)
        for rep in range(2):                                                       # twice: events and buffers are reused
            outs = [torch.full((n,), 7.0, dtype=dtype).pin_memory() for _ in modes]
            got = hc.run_multi(h_in, [(o, m, s, sd) for o, (m, s, sd) in zip(outs, modes)])
            for o, (m, s, sd) in zip(got, modes):
                want = F.cast_amax_scaled(x, m, seed=sd) if s == "amax" else F.cast(x, m, scale=s, seed=sd)
                assert same(o, want.cpu()), (m, s, chunk, rep)
    if dtype == torch.float32:
        O.set_nan_mode(False)
        xn = h_in.numpy()
        s = O.scale_from_amax(O.amax(xn), 448.0)
        assert O.same_bits_or_both_nan(outs[4].numpy(), O.cast(xn, "E4M3_RNE", scale=s)).all()


def test_host_multi_argument_errors(F):
    L = F._lib
    hc = F.HostCaster(torch.float32, chunk_elems=1 << 12)
    h = torch.zeros(100).pin_memory()
    with pytest.raises(ValueError):
        hc.run_multi(h, [(torch.zeros(100).pin_memory(), "BFLOAT16_RNE", "amax", 0)])      # BFLOAT16 takes no scale
    with pytest.raises(ValueError):
        hc.run_multi(h, [(torch.zeros(100).pin_memory(), "E4M3_RNE", 70000.0, 0)])
    with pytest.raises(ValueError):
        hc.run_multi(h, [(torch.zeros(50).pin_memory(), "E4M3_RNE", 1.0, 0)])
    assert hc.run_multi(torch.zeros(0), [(torch.zeros(0), "E4M3_RNE", 1.0, 0)]) is not None
