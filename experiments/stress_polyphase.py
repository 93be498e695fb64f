"""Stress: the SNAC polyphase WNConvTranspose1d case that failed once inside the full GPU suite - repeat it with the allocator
in different states and report any run that differs from the first."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from mlx_audio_b200.codec.models.snac import layers as S

torch.manual_seed(0)
bad = 0
for stride, cin, cout in [(7, 96, 64), (8, 64, 32), (3, 40, 20), (2, 32, 32)]:
    r = np.random.Generator(np.random.PCG64(stride))
    m = S.WNConvTranspose1d(cin, cout, 2 * stride, stride=stride, padding=-(-stride // 2), output_padding=stride % 2)
    v = r.standard_normal((cin, 2 * stride, cout), dtype=np.float32)
    gq = r.uniform(0.5, 1.5, (cin, 1, 1)).astype(np.float32)
    b = (0.1 * r.standard_normal(cout, dtype=np.float32)).astype(np.float32)
    m.load_weights({"weight_v": v, "weight_g": gq, "bias": b})
    x = torch.from_numpy(r.standard_normal((2, 37, cin), dtype=np.float32)).cuda()
    first = None
    for it in range(300):
        # dirty the allocator: blocks of assorted sizes filled with NaN, freed again
        junk = [torch.full((int(n),), float("nan"), device="cuda") for n in np.random.randint(1000, 300000, size=6)]
        del junk
        if it % 50 == 0:
            m._folded = None            # re-fold / re-pack the weights into fresh (dirty) memory
        y = m(x)
        torch.cuda.synchronize()
        if first is None:
            first = y.clone()
        elif not torch.equal(first, y):
            d = (first - y)
            nanc = int(torch.isnan(y).sum())
            idx = torch.nonzero(~(first == y))
            print(f"stride {stride}: run {it} differs: max |d| {float(d.abs().nan_to_num(1e9).max()):.3e}, NaNs {nanc}, "
                  f"{idx.shape[0]} elements, first at {idx[0].tolist()} last at {idx[-1].tolist()}")
            bad += 1
            if bad > 6:
                sys.exit(1)
print("differing runs:", bad)
