"""Parity gates of the bf16 tensor-core mode, in ONE place, with the numbers they were derived from.

fp32 mode: the north star's gates (waveform max-abs <= 1e-3, log-mel L1 <= 1e-3) are written in the tests themselves.
bf16 mode (bf16 operands, fp32 accumulate): the gates are ~3x the errors MEASURED on B200 against the oracle
(`record()` appends every measurement to gpurun_out/parity_measured.jsonl; the table in DESIGN.md section 2 is
copied from it), so that a regression of a few x fails instead of hiding under a 10 %-of-peak bound.
"""
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# measured on B200 (round 2, profiles/r02_parity_measured.jsonl): see DESIGN.md section 2 for the per-case table
BF16_MAX_ABS = 3e-2      # waveform max-abs error (signal peak ~0.3)
BF16_REL_L2 = 5e-2       # ||y - ref|| / ||ref||
BF16_MEL_L1 = 5e-2       # mean |log-mel-80(y) - log-mel-80(ref)|


def record(case: str, **vals) -> None:
    """Append one measurement (the numbers the gates are derived from) to gpurun_out/parity_measured.jsonl."""
    try:
        d = os.path.join(ROOT, "gpurun_out")
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, "parity_measured.jsonl"), "a") as f:
            f.write(json.dumps({"case": case, **{k: float(v) for k, v in vals.items()}}) + "\n")
    except OSError:
        pass
    print("parity", case, {k: float(v) for k, v in vals.items()})
