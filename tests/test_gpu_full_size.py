"""BASELINE-size checks (configs[0]: 16 Mi fp32 E4M3_RNE vs the reference CPU op; configs[1]: up to 4 Gi
elements) through size-independent properties: idempotence, monotonicity, sign symmetry, slab invariance (also for
stochastic rounding), half-ulp error bound, and oracle spot checks beyond the 2^31 / 2^32 element marks
(64-bit indexing)."""
import numpy as np
import pytest
import torch

from oracle import fpemu_oracle as O
from oracle import ref_loader

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def F():
    import fpemu_b200
    fpemu_b200._lib.lib()
    return fpemu_b200


def test_config0_16Mi_e4m3_rne_equals_reference_cpu_op_and_oracle(F):
    """BASELINE configs[0]: x = randn(16 Mi, seed 0), E4M3_RNE, scale 1 — bit-exact against the reference's own CPU
    op (oracle/_ref build, when it runs on this host) and against the C oracle."""
    n = 16 * 1024 * 1024
    x = torch.randn(n, dtype=torch.float32, generator=torch.Generator().manual_seed(0))
    got = F.cast(x.cuda(), "E4M3_RNE").cpu()
    want = torch.from_numpy(O.cast(x.numpy(), "E4M3_RNE"))
    assert torch.equal(got.view(torch.int32), want.view(torch.int32))
    cpu, label = ref_loader.best_cpu()
    if cpu is not None:
        ref = cpu.forward(x, "E4M3_RNE", n, False, 1.0, False, 1)[0]
        assert torch.equal(got.view(torch.int32), ref.view(torch.int32)), "differs from the reference CPU op (%s build)" % label


@pytest.mark.parametrize("mode,mant", [("E5M2_RNE", 2), ("E4M3_RNE", 3), ("E3M4_RNE", 4), ("BFLOAT16_RNE", 7)])
def test_properties_at_2_pow_30(F, mode, mant):
    n = 1 << 30
    g = torch.Generator(device="cuda").manual_seed(7)
    x = torch.randn(n, device="cuda", generator=g)
    q = F.cast(x, mode)
    # idempotent: a quantized tensor is a fixed point
    assert torch.equal(F.cast(q, mode).view(torch.int32), q.view(torch.int32))
    # sign symmetry (RNE is symmetric; zeros keep their sign except the documented +0 flush of BFLOAT16)
    qn = F.cast(-x, mode)
    assert torch.equal(qn, -q)
    # in-place == out-of-place at full size
    y = x.clone()
    F.cast(y, mode, out=y)
    assert torch.equal(y.view(torch.int32), q.view(torch.int32))
    del y, qn
    # half-ulp bound for values well inside the format's normal range: |q - h| <= 2^-(mant+1) * 2^floor(log2|h|)
    sl = slice(0, 1 << 26)
    h = x[sl].half().float() if not mode.startswith("BFLOAT16") else x[sl]
    inside = (h.abs() > 0.3) & (h.abs() < 16)
    err = (q[sl] - h).abs()[inside]
    bound = torch.exp2(torch.floor(torch.log2(h.abs()[inside])) - (mant + 1))
    assert bool((err <= bound).all())
    # monotone: sorted input gives sorted output
    xs, _ = torch.sort(x[: 1 << 27])
    qs = F.cast(xs, mode)
    assert bool((qs[1:] >= qs[:-1]).all())


@pytest.mark.parametrize("mode", ["E4M3_RNE", "E5M2_STOCHASTIC", "BFLOAT16_STOCHASTIC"])
def test_slab_invariance_at_2_pow_30(F, mode):
    """Casting 8 row slabs with their element offsets == casting the whole tensor (what the 8-GPU sharding relies on)."""
    n = 1 << 30
    x = torch.randn(n, device="cuda", generator=torch.Generator(device="cuda").manual_seed(3))
    whole = F.cast(x, mode, seed=42)
    out = torch.empty_like(x)
    step = n // 8
    for r in range(8):
        F.cast(x[r * step:(r + 1) * step], mode, out=out[r * step:(r + 1) * step], seed=42, elem_offset=r * step)
    assert torch.equal(out.view(torch.int32), whole.view(torch.int32))


@pytest.mark.parametrize("mode", ["E4M3_RNE", "E3M4_STOCHASTIC"])
def test_beyond_4Gi_elements_64bit_indexing(F, mode):
    """2^32 + 17 bf16 elements (8 GiB in, 8 GiB out): oracle spot checks at the start, around 2^31, around 2^32 and
    on the ragged tail."""
    n = (1 << 32) + 17
    free, _ = torch.cuda.mem_get_info()
    if free < 20 * (1 << 30):
        pytest.skip("not enough free device memory")
    x = torch.empty(n, dtype=torch.bfloat16, device="cuda")
    chunk = 1 << 28
    g = torch.Generator(device="cuda").manual_seed(11)
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        x[s:e] = (torch.randn(e - s, device="cuda", generator=g) * 3).to(torch.bfloat16)
    q = F.cast(x, mode, seed=5)
    for start in (0, (1 << 31) - 1000, (1 << 32) - 1000, n - 2000):
        stop = min(n, start + 2000)
        xs = x[start:stop].view(torch.int16).cpu().numpy().view(np.uint16)
        want = O.cast(xs, mode, seed=5, elem_offset=start)
        got = q[start:stop].view(torch.int16).cpu().numpy().view(np.uint16)
        assert O.same_bits_or_both_nan(got, want).all(), "window at %d" % start
    a = F.amax(x).item()
    assert a == x[: 1 << 28].float().abs().max().item() or a >= x[: 1 << 28].float().abs().max().item()
    # the 8-bit wire form over the same tensor: codes and their decoded values in the same windows
    del q
    codes = F.encode(x, mode, seed=5)
    back = F.decode(codes, mode.split("_")[0], dtype=torch.bfloat16)
    for start in (0, (1 << 31) - 1000, (1 << 32) - 1000, n - 2000):
        stop = min(n, start + 2000)
        xs = x[start:stop].view(torch.int16).cpu().numpy().view(np.uint16)
        want = O.encode(xs, mode, seed=5, elem_offset=start)
        assert (codes[start:stop].cpu().numpy() == want).all(), "codes, window at %d" % start
        got = back[start:stop].view(torch.int16).cpu().numpy().view(np.uint16)
        assert (got == O.decode(want, mode.split("_")[0], dtype=np.uint16)).all(), "decode, window at %d" % start
