"""Seeded random configurations of the cast and the 8-bit encode against the C oracle: mode x dtype x size (clustered
around CTA and vector boundaries) x scale (host / device scalar) x pointer offset x element offset x in-place.
Bit-exact, NaN compared as NaN.  A failure message carries the configuration."""
import numpy as np
import pytest
import torch

from oracle import fpemu_oracle as O
from test_gpu_cast_parity import assert_same, wide
from test_gpu_codec import as_dtype, codes_equal

pytestmark = pytest.mark.gpu

CAST_MODES = ["E5M2_RNE", "E5M2_STOCHASTIC", "E5M2_RTZ", "E5M2_RNAZ", "E5M2_RNTZ", "E5M2_RPINF", "E5M2_RNINF",
              "E5M2_DAZ_RNE", "E5M2_DAZ_STOCHASTIC", "E5M2_DAZ_RNAZ", "E4M3_RNE", "E4M3_STOCHASTIC", "E4M3_RTZ",
              "E3M4_RNE", "E3M4_STOCHASTIC", "E3M4_RTZ", "BFLOAT16_RNE", "BFLOAT16_STOCHASTIC", "BFLOAT16_RTZ",
              "FLOAT16_RNE", "FLOAT16_STOCHASTIC", "E4M3_IEEE_RNE", "E4M3_IEEE_STOCHASTIC", "FP4_NEAREST"]
CODE_MODES = [m for m in CAST_MODES if m.split("_")[0] in ("E5M2", "E4M3", "E3M4") and "IEEE" not in m]
BOUNDARIES = [8, 16, 256, 2048, 4096, 8192, 16384, 32768]          # vector, warp and CTA tile sizes of the kernels


@pytest.fixture(scope="module")
def F():
    import fpemu_b200
    fpemu_b200._lib.lib()
    return fpemu_b200


@pytest.fixture(autouse=True)
def _cuda_nan_mode():
    O.set_nan_mode(False)
    yield


def random_config(rng, modes):
    mode = modes[rng.integers(len(modes))]
    dt = ["f32", "f16", "bf16"][rng.integers(3)]
    if rng.random() < 0.6:
        n = int(BOUNDARIES[rng.integers(len(BOUNDARIES))] * rng.integers(1, 4) + rng.integers(-3, 4))
    else:
        n = int(rng.integers(1, 70000))
    n = max(n, 1)
    scale = [1.0, 1.0, 0.125, 3.7, 94.09, 448.0 / 3.1][rng.integers(6)]
    if mode.split("_")[0] in ("BFLOAT16", "FLOAT16"):
        scale = 1.0
    ptr_off = int([0, 0, 0, 1, 3, 8, 16][rng.integers(7)])            # elements skipped at the front of the buffer
    elem_off = int([0, 0, 16, 48, 5, 1 << 33][rng.integers(6)])       # global index of the first element
    return dict(mode=mode, dt=dt, n=n, scale=scale, ptr_off=ptr_off, elem_off=elem_off,
                device_scale=bool(rng.random() < 0.3) and scale != 1.0, inplace=bool(rng.random() < 0.3),
                seed=int(rng.integers(1, 1 << 40)), data_seed=int(rng.integers(1 << 30)))


def bits32(a, dt):
    if dt == "bf16":
        return (a.astype(np.uint32) << 16).view(np.float32)
    return a.astype(np.float32) if dt == "f16" else a


@pytest.mark.parametrize("chunk", range(6))
def test_random_cast_configurations(F, chunk):
    rng = np.random.default_rng(1000 + chunk)
    for _ in range(40):
        c = random_config(rng, CAST_MODES)
        xn, xt = as_dtype(wide(c["n"] + c["ptr_off"], c["data_seed"]), c["dt"])
        xn, xt = xn[c["ptr_off"]:], xt[c["ptr_off"]:]
        want = O.cast(xn, c["mode"], scale=c["scale"], seed=c["seed"], elem_offset=c["elem_off"])
        kw = dict(seed=c["seed"], elem_offset=c["elem_off"])
        if c["device_scale"]:
            kw["d_scale"] = torch.tensor([c["scale"]], device="cuda")
        else:
            kw["scale"] = c["scale"]
        if c["inplace"]:
            kw["out"] = xt
        got = F.cast(xt, c["mode"], **kw)
        if xt.dtype == torch.bfloat16:
            got_np = got.view(torch.int16).cpu().numpy().view(np.uint16)
        else:
            got_np = got.cpu().numpy()
        assert_same(bits32(got_np, c["dt"]), bits32(want, c["dt"]), str(c))


@pytest.mark.parametrize("chunk", range(4))
def test_random_encode_decode_configurations(F, chunk):
    rng = np.random.default_rng(2000 + chunk)
    for _ in range(40):
        c = random_config(rng, CODE_MODES)
        fmt = c["mode"].split("_")[0]
        xn, xt = as_dtype(wide(c["n"] + c["ptr_off"], c["data_seed"]), c["dt"])
        xn, xt = xn[c["ptr_off"]:], xt[c["ptr_off"]:]
        kw = dict(seed=c["seed"], elem_offset=c["elem_off"])
        if c["device_scale"]:
            kw["d_scale"] = torch.tensor([c["scale"]], device="cuda")
        else:
            kw["scale"] = c["scale"]
        buf = torch.zeros(c["n"] + 64, dtype=torch.uint8, device="cuda")
        codes = F.encode(xt, c["mode"], out=buf[c["ptr_off"]:c["ptr_off"] + c["n"]], **kw)
        want = O.encode(xn, c["mode"], scale=c["scale"], seed=c["seed"], elem_offset=c["elem_off"])
        codes_equal(codes, want, str(c))
        assert int(buf[:c["ptr_off"]].sum()) == 0 and int(buf[c["ptr_off"] + c["n"]:].sum()) == 0, str(c)
        np_dt = {"f32": np.float32, "f16": np.float16, "bf16": np.uint16}[c["dt"]]
        back = F.decode(codes, fmt, dtype=xt.dtype, scale=c["scale"])
        back_np = back.view(torch.int16).cpu().numpy().view(np.uint16) if xt.dtype == torch.bfloat16 else back.cpu().numpy()
        assert_same(bits32(back_np, c["dt"]), bits32(O.decode(want, fmt, dtype=np_dt, scale=c["scale"]), c["dt"]), str(c))


@pytest.mark.parametrize("chunk", range(4))
def test_input_and_output_misaligned_by_different_amounts(F, chunk):
    """A caller-supplied output whose offset from a 32-byte boundary differs from the input's (cast_shift_kernel: aligned
    stores, input assembled from two aligned blocks), any element offset, guard zones around the output."""
    rng = np.random.default_rng(3000 + chunk)
    modes = ["E5M2_RNE", "E5M2_STOCHASTIC", "E5M2_RTZ", "E4M3_RNE", "E4M3_STOCHASTIC", "E4M3_RTZ", "E3M4_RNE",
             "E3M4_STOCHASTIC", "BFLOAT16_RNE", "BFLOAT16_STOCHASTIC", "BFLOAT16_RTZ", "E4M3_IEEE_RNE", "FLOAT16_STOCHASTIC"]
    for _ in range(40):
        c = random_config(rng, modes)
        in_off, out_off = int(rng.integers(0, 16)), int(rng.integers(0, 16))
        xn, xt = as_dtype(wide(c["n"] + in_off, c["data_seed"]), c["dt"])
        xn, xt = xn[in_off:], xt[in_off:]
        guard = 40
        big = torch.full((guard + out_off + c["n"] + guard,), 77.0, dtype=xt.dtype, device="cuda")
        out = big[guard + out_off: guard + out_off + c["n"]]
        want = O.cast(xn, c["mode"], scale=c["scale"], seed=c["seed"], elem_offset=c["elem_off"])
        F.cast(xt, c["mode"], out=out, scale=c["scale"], seed=c["seed"], elem_offset=c["elem_off"])
        got_np = out.view(torch.int16).cpu().numpy().view(np.uint16) if xt.dtype == torch.bfloat16 else out.cpu().numpy()
        what = "%s in_off=%d out_off=%d" % (c, in_off, out_off)
        assert_same(bits32(got_np, c["dt"]), bits32(want, c["dt"]), what)
        assert bool((big[:guard + out_off] == 77.0).all()) and bool((big[guard + out_off + c["n"]:] == 77.0).all()), what
