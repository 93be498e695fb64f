"""Where does conv1d_tc_persist_kernel spend its time on the memory-bound generator layers?  Runs the noise_convs[1]
(1x1, 24->128 at 120T+1) and ups[1] (polyphase 256->128 x6) shapes with the epilogue outputs switched on and off."""
import ctypes as C, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mlx_audio_b200 import _lib as L
from mlx_audio_b200._lib import ConvDesc
lib = L.load()
bf = torch.bfloat16

def run(name, B, L_in, C_in, x_pitch, taps, N, C_out, up, up_off, M, L_true, L_out, left_pad, pad_left, res, out, act, stats):
    x = torch.randn(B, L_in, x_pitch, device="cuda").to(bf)
    Cw = (C_in + 63) // 64 * 64
    w = (torch.randn(N, taps * Cw, device="cuda") / (C_in * taps) ** 0.5).to(bf)
    bias = torch.zeros(C_out, device="cuda")
    d = ConvDesc()
    d.B, d.L_in, d.C_in, d.x_pitch = B, L_in, C_in, x_pitch
    d.taps, d.dil, d.pad_left, d.stride, d.Cw = taps, 1, pad_left, 1, Cw
    d.M, d.N, d.C_out, d.up, d.up_off, d.L_true, d.left_pad, d.L_out = M, N, C_out, up, up_off, L_true, left_pad, L_out
    d.bias = bias.data_ptr(); d.scale = 1.0
    keep = []
    if res:
        r = torch.randn(B, L_out, C_out, device="cuda"); keep.append(r)
        d.res, d.res_pitch, d.res_shift, d.L_res = r.data_ptr(), C_out, 0, L_out
    if out:
        o = torch.empty(B, L_out, C_out, device="cuda"); keep.append(o)
        d.out, d.out_pitch = o.data_ptr(), C_out
    if act:
        a = torch.empty(B, L_out, C_out, device="cuda", dtype=bf); keep.append(a)
        d.out_act, d.act_pitch, d.act_is_bf16, d.act_slope = a.data_ptr(), C_out, 1, 1.0
    if stats:
        s = torch.zeros(B, C_out, 2, dtype=torch.float64, device="cuda"); keep.append(s)
        d.stats = s.data_ptr()
    for _ in range(2):
        L.check(lib.kk_conv1d_bf16_tc(x.data_ptr(), w.data_ptr(), C.byref(d), 0, L.stream()))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        L.check(lib.kk_conv1d_bf16_tc(x.data_ptr(), w.data_ptr(), C.byref(d), 0, L.stream()))
    e1.record(); torch.cuda.synchronize()
    print(f"{name}: res={int(res)} out={int(out)} act={int(act)} stats={int(stats)}: {e0.elapsed_time(e1) / 3:.3f} ms")

B, L3, L20 = 64, 38401, 6400
for res, out, act, stats in [(0, 0, 1, 1), (0, 0, 1, 0), (0, 1, 0, 0), (0, 1, 0, 1), (0, 1, 1, 1)]:
    run("noise1 1x1 24->128", B, 6 * (L20 + 1), 24, 24, 1, 128, 128, 1, 0, L3, L3, L3, 0, 0, res, out, act, stats)
for res, out, act, stats in [(1, 0, 1, 1), (1, 0, 1, 0), (0, 0, 1, 0), (1, 1, 0, 1), (1, 1, 0, 0), (0, 1, 0, 0)]:
    run("ups1 256->6x128", B, L20, 256, 256, 2, 768, 128, 6, 3, L20 + 1, L20 * 6, L3, 1, 1, res, out, act, stats)
