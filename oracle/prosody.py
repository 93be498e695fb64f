"""CPU oracle (TEST INFRASTRUCTURE ONLY - never imported by the product path) for ``ProsodyPredictor.F0Ntrain``
(mlx_audio/tts/models/kokoro/modules.py:355-377): the shared bi-LSTM (modules.py:93-286) followed by two stacks of
three AdainResBlk1d (modules.py:299-337; istftnet.py:825-899) and 1x1 Conv1d projections (modules.py:339-340).
float32 torch-CPU restatement in the reference's operation order; pinned against the reference's own source executed
on oracle/mlx_shim (tests/golden/prosody_b2_l12.npz, tests/test_oracle.py)."""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import numpy as np
import torch

from .istftnet import DecoderOracle


class F0NOracle:
    def __init__(self, weights: Dict[str, np.ndarray]):
        w = {(k[len("predictor."):] if k.startswith("predictor.") else k): v for k, v in weights.items()}
        # DecoderOracle only supplies the block helpers (_adain_resblk1d, _conv_weighted, _adain) over its weight dict
        self.blocks = DecoderOracle(w)
        self.w = self.blocks.w

    def lstm(self, x_blc: torch.Tensor) -> torch.Tensor:
        """LSTM.__call__ (modules.py:239-286): both directions, zero initial state, gates i, f, g, o (:178-190)."""
        outs = []
        for d in ("forward", "backward"):
            Wx, Wh = self.w[f"shared.Wx_{d}"], self.w[f"shared.Wh_{d}"]
            xp = (self.w[f"shared.bias_ih_{d}"] + self.w[f"shared.bias_hh_{d}"]) + x_blc @ Wx.T      # addmm :156
            B, Lq, _ = x_blc.shape
            H = Wh.shape[1]
            h = torch.zeros(B, H)
            c = torch.zeros(B, H)
            hs = [None] * Lq
            order = range(Lq) if d == "forward" else range(Lq - 1, -1, -1)
            for t in order:
                i, f, g, o = torch.chunk(xp[:, t] + h @ Wh.T, 4, dim=-1)
                c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(g)
                h = torch.sigmoid(o) * torch.tanh(c)
                hs[t] = h
            outs.append(torch.stack(hs, dim=1))
        return torch.cat(outs, dim=-1)                                                                # :279

    def heads(self, hidden_blc: torch.Tensor, s: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        res = []
        for head in ("F0", "N"):
            y = hidden_blc.transpose(1, 2)                                                            # :361 / :370
            y = self.blocks._adain_resblk1d(f"{head}.0", y, s)
            y = self.blocks._adain_resblk1d(f"{head}.1", y, s, upsample=True)
            y = self.blocks._adain_resblk1d(f"{head}.2", y, s)
            pw, pb = self.w[f"{head}_proj.weight"], self.w[f"{head}_proj.bias"]                        # (1, 1, C), (1,)
            res.append(torch.einsum("bcl,c->bl", y, pw[0, 0]) + pb[0])                                 # :366 / :374
        return res[0], res[1]

    def __call__(self, x=None, s=None, hidden: Optional[np.ndarray] = None):
        """x (B, d_hid + style, L), s (B, style) -> (F0, N) each (B, 2L) float32.  ``hidden`` (B, L, d_hid) skips the
        LSTM.  Also returns nothing else; use ``hidden_of`` for the LSTM output."""
        with torch.no_grad():
            s_t = torch.from_numpy(np.asarray(s, np.float32))
            h = self.hidden_of(x, as_torch=True) if hidden is None else torch.from_numpy(np.asarray(hidden, np.float32))
            F0, N = self.heads(h, s_t)
        return F0.numpy(), N.numpy()

    def hidden_of(self, x, as_torch: bool = False):
        with torch.no_grad():
            h = self.lstm(torch.from_numpy(np.asarray(x, np.float32)).transpose(1, 2))
        return h if as_torch else h.numpy()
