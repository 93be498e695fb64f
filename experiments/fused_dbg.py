"""Debug driver: where does conv1d_adain_tc_kernel wait?  Runs single fused convs at bench size on the lab build
(`KK_TUNE=1 KK_FUSED_DBG=1 python -m mlx_audio_b200._build` -> mlx_audio_b200/libkokoro_b200_lab.so) and prints per-role
barrier wait time as a share of kernel time."""
import ctypes as C, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mlx_audio_b200 import _lib as L
L.LIB_PATH = os.path.join(ROOT, "mlx_audio_b200", "libkokoro_b200_lab.so")
from mlx_audio_b200._lib import AdainIn, ConvDesc
lib = L.load()
lib.kk_debug_fused_buffer.argtypes = [C.c_void_p]; lib.kk_debug_fused_buffer.restype = C.c_int
NAMES = ["Bprod:b_empty", "MMA:tmem_empty", "MMA:a_full", "MMA:b_full", "Xprod:empty", "XF:data", "XF:slot_empty",
         "EPI:tmem_full", "EPI:res_full", "-", "MMA:total"]

def run(B, Lr, Cc, k, dil, x_bf16, with_res, with_out, with_stats=True):
    dev = "cuda"
    x = torch.randn(B, Lr, Cc, device=dev)
    xd = x.to(torch.bfloat16).contiguous() if x_bf16 else x
    stats = torch.stack([x.double().sum(1), (x.double() ** 2).sum(1)], -1).contiguous()
    h = torch.randn(B, 2 * Cc, device=dev) * 0.1
    alpha = torch.rand(Cc, device=dev) + 0.5
    w = (torch.randn(Cc, k * Cc, device=dev) / (Cc * k) ** 0.5).to(torch.bfloat16)
    bias = torch.zeros(Cc, device=dev)
    res = torch.randn(B, Lr, Cc, device=dev) if with_res else None
    if with_res == "bf16": res = res.to(torch.bfloat16)
    out = torch.empty(B, Lr, Cc, device=dev) if with_out else None
    oact = None if with_out else torch.empty(B, Lr, Cc, device=dev, dtype=torch.bfloat16)
    ostats = torch.zeros(B, Cc, 2, dtype=torch.float64, device=dev)
    dbg = torch.zeros(148 * 16, dtype=torch.int64, device=dev)
    L.check(lib.kk_debug_fused_buffer(dbg.data_ptr()))
    a = AdainIn(); a.x, a.x_is_bf16, a.x_pitch = xd.data_ptr(), int(x_bf16), Cc
    a.stats, a.h, a.h_stride, a.alpha, a.act, a.slope = stats.data_ptr(), h.data_ptr(), 2 * Cc, alpha.data_ptr(), 3, 0.0
    d = ConvDesc(); d.B, d.L_in, d.C_in, d.x_pitch = B, Lr, Cc, Cc
    d.taps, d.dil, d.pad_left, d.stride, d.Cw = k, dil, (k * dil - dil) // 2, 1, Cc
    d.M, d.N, d.C_out, d.up, d.up_off, d.L_true, d.left_pad, d.L_out = Lr, Cc, Cc, 1, 0, Lr, 0, Lr
    d.bias = bias.data_ptr()
    if with_res: d.res, d.res_pitch, d.res_shift, d.L_res = res.data_ptr(), Cc, 0, Lr
    d.res_is_bf16 = int(with_res == "bf16")
    d.scale = 1.0
    if with_out: d.out, d.out_pitch = out.data_ptr(), Cc
    else: d.out_act, d.act_pitch, d.act_is_bf16, d.act_slope = oact.data_ptr(), Cc, 1, 1.0
    if with_stats: d.stats = ostats.data_ptr()
    for _ in range(2):
        L.check(lib.kk_conv1d_adain_bf16_tc(C.byref(a), w.data_ptr(), C.byref(d), L.stream()))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); L.check(lib.kk_conv1d_adain_bf16_tc(C.byref(a), w.data_ptr(), C.byref(d), L.stream())); e1.record()
    torch.cuda.synchronize()
    v = dbg.view(148, 16).double().mean(0).cpu().numpy()
    tot = v[10]
    print(f"C={Cc} k={k} d={dil} x_bf16={x_bf16} res={with_res} out={'f32' if with_out else 'bf16'}: {e0.elapsed_time(e1):.3f} ms; "
          + "  ".join(f"{n}={v[i] / tot * 100:.0f}%" for i, n in enumerate(NAMES) if n != "-" and i != 10))

def run_post(B, Lr):
    """conv_post shape: C=128 -> 22, k=7, bf16 input that is already activated (no normalisation, no activation)."""
    dev = "cuda"
    x = torch.randn(B, Lr, 128, device=dev).to(torch.bfloat16).contiguous()
    w = (torch.randn(22, 7 * 128, device=dev) / 30).to(torch.bfloat16)
    bias = torch.zeros(24, device=dev)
    out = torch.empty(B, Lr, 24, device=dev)
    dbg = torch.zeros(148 * 16, dtype=torch.int64, device=dev)
    L.check(lib.kk_debug_fused_buffer(dbg.data_ptr()))
    a = AdainIn(); a.x, a.x_is_bf16, a.x_pitch, a.act = x.data_ptr(), 1, 128, 0
    d = ConvDesc(); d.B, d.L_in, d.C_in, d.x_pitch = B, Lr, 128, 128
    d.taps, d.dil, d.pad_left, d.stride, d.Cw = 7, 1, 3, 1, 128
    d.M, d.N, d.C_out, d.up, d.up_off, d.L_true, d.left_pad, d.L_out = Lr, 22, 22, 1, 0, Lr, 0, Lr
    d.bias, d.scale, d.out, d.out_pitch = bias.data_ptr(), 1.0, out.data_ptr(), 24
    for _ in range(2):
        L.check(lib.kk_conv1d_adain_bf16_tc(C.byref(a), w.data_ptr(), C.byref(d), L.stream()))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); L.check(lib.kk_conv1d_adain_bf16_tc(C.byref(a), w.data_ptr(), C.byref(d), L.stream())); e1.record()
    torch.cuda.synchronize()
    v = dbg.view(148, 16).double().mean(0).cpu().numpy()
    print(f"conv_post 128->22 k=7: {e0.elapsed_time(e1):.3f} ms; total {v[10]:.0f} clk; "
          + "  ".join(f"{n}={v[i] / v[10] * 100:.0f}%" for i, n in enumerate(NAMES) if n != "-" and i != 10))


import os
if os.environ.get("DBG_SET") == "post":
    run_post(64, 38401)
    sys.exit(0)
if os.environ.get("DBG_SET") == "res":   # KK_L2PF=5 / 21 / 37: residual path whole / fetched not added / added not fetched
    for k in (3, 7):
        run(64, 38401, 128, k, 1, True, False, False)
        run(64, 38401, 128, k, 1, True, "bf16", False)
    sys.exit(0)
if os.environ.get("DBG_SET") == "nostats":   # how much of an epilogue-bound launch is the statistics arithmetic?
    for k in (3, 7):
        run(64, 38401, 128, k, 1, True, "bf16", False)
        run(64, 38401, 128, k, 1, True, "bf16", False, with_stats=False)
    sys.exit(0)
if os.environ.get("DBG_SET") == "s16":
    for k in (3, 7, 11):
        run(64, 38401, 128, k, 1, False, False, False)   # fp32 in -> bf16 out (fp32-stream conv1)
        run(64, 38401, 128, k, 1, True, False, False)    # bf16 in -> bf16 out (bf16-stream conv1)
        run(64, 38401, 128, k, 1, True, "bf16", False)   # bf16 in + bf16 residual -> bf16 out (bf16-stream conv2)
        run(64, 38401, 128, k, 1, True, True, True)      # bf16 in + fp32 residual -> fp32 out (fp32-stream conv2)
    for k in (3, 7):
        run(64, 6400, 256, k, 1, False, False, False)
        run(64, 6400, 256, k, 1, True, False, False)
        run(64, 6400, 256, k, 1, True, "bf16", False)
        run(64, 6400, 256, k, 1, True, True, True)
    sys.exit(0)
if os.environ.get("DBG_SET") == "c256":
    run(64, 6400, 256, 11, 1, False, False, False)   # fp32 in, bf16 out
    run(64, 6400, 256, 11, 1, True, False, False)    # bf16 in, bf16 out
    run(64, 6400, 256, 11, 1, True, False, True)     # bf16 in, f32 out (TMA store), no residual
    run(64, 6400, 256, 11, 1, True, True, True)      # bf16 in, residual, f32 out
    run(64, 6400, 256, 11, 1, False, False, True)    # fp32 in, f32 out
    sys.exit(0)
for k in (3, 11):
    run(64, 38401, 128, k, 1, False, False, False)   # conv1-type
    run(64, 38401, 128, k, 1, True, True, True)      # conv2-type
for k in (3, 7, 11):
    run(64, 6400, 256, k, 1, False, False, False)
    run(64, 6400, 256, k, 1, True, True, True)
run(64, 38401, 128, 7, 1, False, False, False)
run(64, 38401, 128, 7, 1, True, True, True)
