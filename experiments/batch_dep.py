import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from mlx_audio_b200 import synth, utils
from mlx_audio_b200.tts.models.kokoro import build_decoder
B, T = int(os.environ.get("BB", 64)), int(os.environ.get("TT", 320))
dec = build_decoder(synth.synth_decoder_weights(seed=2000), dtype="bf16")
inp = synth.synth_decoder_inputs(B, T, seed=1000, with_noise=False)
noise = torch.randn(1, 600 * T, 9, device="cuda", generator=torch.Generator(device="cuda").manual_seed(5))
nb = torch.zeros(B, 600 * T, 9, device="cuda"); r = min(17, B - 1); nb[r] = noise[0]
yb = dec(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], noise=nb, capture_taps=True).clone()
tb = {k: v[r:r+1].clone() for k, v in dec.taps(B, T).items()}
y1 = dec(inp["asr"][r:r+1], inp["F0_curve"][r:r+1], inp["N"][r:r+1], inp["s"][r:r+1], noise=noise, capture_taps=True).clone()
t1 = dec.taps(1, T)
for k in tb:
    a, b = tb[k].float(), t1[k].float()
    print(f"{k:14s} rel {float((a-b).norm()/(b.norm()+1e-30)):.3e} max {float((a-b).abs().max()):.3e}")
print("audio rel", float((yb[r]-y1[0]).norm()/y1[0].norm()))
x = 0.1 * torch.randn(4, 720000, device="cuda", generator=torch.Generator(device="cuda").manual_seed(1))
S = utils.stft(x, n_fft=1024, hop_length=256); S1 = utils.stft(x[2], n_fft=1024, hop_length=256)
print("stft batch vs single max abs", float((S[2]-S1).abs().max()))
