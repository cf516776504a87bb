import sys, numpy as np, torch
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/fp8-emulation-toolkit_b200')
import fpemu_b200 as F
from oracle import fpemu_oracle as O
rng=np.random.default_rng(21); n=1<<15
x = rng.standard_normal(n).astype(np.float32) * np.exp2(rng.integers(-30, 20, n)).astype(np.float32)
spec = np.array([0.0, -0.0, np.inf, -np.inf, np.nan, -np.nan, 1e-40, -1e-40, 65504, 65520, -65520, 3.4e38, 1e-45,448, 464, 480, -448, 30, 31, -31, 57344, 61440, -61440], dtype=np.float32); x[:len(spec)]=spec
got=F.cast(torch.from_numpy(x).cuda(),"BFLOAT16_STOCHASTIC",seed=0xC0FFEE1234).cpu().numpy()
want=O.cast(x,"BFLOAT16_STOCHASTIC",seed=0xC0FFEE1234)
bad=np.nonzero(~O.same_bits_or_both_nan(got,want))[0]
print(len(bad), 'mismatches')
r=O.rand_bits(0xC0FFEE1234,0,n,3)
for i in bad[:12]: print(i, hex(x.view(np.uint32)[i]), hex(got.view(np.uint32)[i]), hex(want.view(np.uint32)[i]), hex(r[i]))
