"""Parity criteria shared by the -m gpu tests, and the achieved-error log.

Every comparison of a CUDA result with the oracle goes through `check()`, which applies THREE criteria
and records the achieved numbers (printed as a table at the end of the pytest run by conftest.py, and
written to gpurun_out/parity_errors.log on the GPU box):

  1. north_star's tolerance: max|a - ref| / max|ref| < 2e-2 per tensor (1e-3 absolute for LSE);
  2. the usual FlashAttention test criterion: the error is at most TWICE the error of a 16-bit-pipeline
     reference against the exact one.  The pipeline reference is the oracle's `round_intermediates`
     build (P after dropout and dS rounded to the storage dtype before their GEMMs, outputs rounded to
     the storage dtype): the roundings any fp16/bf16 tensor-core implementation cannot avoid;
  3. per row (one (token, head) vector): max_c|a - ref| <= 2e-2 * max(max_c|ref| of that row,
     ROW_FLOOR * max|ref|), so that one bad row cannot hide behind the tensor's global maximum.
"""
from __future__ import annotations

import os

import numpy as np

TOL_REL = 2e-2        # BASELINE.json north_star: relative, on out / dq / dk / dv
TOL_LSE = 1e-3        # absolute, on LSE
ROW_FLOOR = 0.05      # a row's own scale is floored at this fraction of the tensor's max|ref|
LP_FACTOR = 2.0       # criterion 2
LP_SLACK = 2e-5       # added to criterion 2's bound (both errors are ratios to max|ref|)

LOG = []              # (label, tensor, rel err, lp-baseline err or None, worst row ratio)


def rel_err(a, ref):
    a = np.asarray(a, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    den = np.abs(ref).max(initial=0.0)
    if den == 0:
        return float(np.abs(a).max(initial=0.0))
    return float(np.abs(a - ref).max(initial=0.0) / den)


def row_ratio(a, ref):
    """max over rows of max_c|a - ref| / max(row max|ref|, ROW_FLOOR * global max|ref|); rows = all but the last axis."""
    a = np.asarray(a, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    if a.size == 0:
        return 0.0
    g = np.abs(ref).max()
    if g == 0:
        return float(np.abs(a).max())
    err = np.abs(a - ref).max(axis=-1)
    den = np.maximum(np.abs(ref).max(axis=-1), ROW_FLOOR * g)
    return float((err / den).max())


def check(label, tensor, got_f32, ref_f32, lp_f32=None, tol=TOL_REL):
    """Applies the three criteria to one tensor; `lp_f32` is the 16-bit-pipeline reference (or None: criterion 2 skipped)."""
    err = rel_err(got_f32, ref_f32)
    base = rel_err(lp_f32, ref_f32) if lp_f32 is not None else None
    rr = row_ratio(got_f32, ref_f32)
    LOG.append((label, tensor, err, base, rr))
    assert err < tol, f"{label} {tensor}: rel err {err:.3e} >= {tol}"
    if base is not None:
        bound = LP_FACTOR * base + LP_SLACK
        assert err <= bound, f"{label} {tensor}: rel err {err:.3e} > 2 x 16-bit-pipeline reference ({base:.3e})"
    assert rr <= tol, f"{label} {tensor}: worst row ratio {rr:.3e} > {tol}"
    return err


def check_lse(label, got, ref, tol=TOL_LSE):
    got = np.asarray(got)
    ref = np.asarray(ref)
    assert np.array_equal(np.isposinf(got), np.isposinf(ref)), f"{label}: +inf rows of LSE differ"
    fin = np.isfinite(ref)
    err = float(np.abs(got[fin] - ref[fin]).max(initial=0))
    LOG.append((label, "lse(abs)", err, None, None))
    assert err < tol, f"{label}: lse err {err:.3e}"
    return err


def summary_lines():
    out = [f"{'case':58s} {'tensor':9s} {'rel err':>10s} {'16-bit ref':>10s} {'worst row':>10s}"]
    for label, tensor, err, base, rr in LOG:
        out.append(f"{label[:58]:58s} {tensor:9s} {err:10.3e} {('%10.3e' % base) if base is not None else '         -'} "
                   f"{('%10.3e' % rr) if rr is not None else '         -'}")
    return out


def write_log(root):
    d = os.path.join(root, "gpurun_out")
    if LOG and os.path.isdir(d):
        with open(os.path.join(d, "parity_errors.log"), "w") as fh:
            fh.write("\n".join(summary_lines()) + "\n")
