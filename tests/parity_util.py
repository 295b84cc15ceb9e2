"""Shared comparison rules of the GPU parity tests (and of bench.py's `parity` key).

Tolerances (BASELINE.json north_star): FP64 GRFs within 1e-6 relative of the exact solution and the
same primal-active set.  Two error measures are used:

* element-wise (`elem_err`):  |got - ref| / max(|ref|, 1e-3 * max(||ref||_inf, 1 N))  per component
  — the north-star assertion (<= REL_TOL on every component, small tangential forces included; the
  1 N floor on the norm keeps an all-zero answer — a robot in free fall — from comparing rounding
  noise of 1e-15 N with exact zeros);
* norm-wise (`norm_err`):     max|got - ref| / max(1, ||ref||_inf)           per problem — the
  regression guard on exactly polished answers: median ~2e-12, 99 % of the problems below
  GUARD_TOL = 1e-9 and none above GUARD_MAX = 1e-7.  (The guard is a quantile because both sides
  solve an equality-constrained system with cond(H) up to ~1e7: 1e7 x 2^-53 ~ 1e-9 is the accuracy
  either solver can deliver on the worst-conditioned problems of a 65 536 batch.)
"""
import numpy as np

REL_TOL = 1e-6     # north_star: FP64 GRFs within 1e-6 relative
GUARD_TOL = 1e-9   # regression guard: 99 % quantile of the norm-wise error of status-0 answers
GUARD_MAX = 1e-7   # ... and its maximum
F32_TOL = 1e-3     # north_star: FP32 variant within 1e-3 relative
BAND_FRAC_MAX = 1e-4  # forgiven "tie" rows (oracle residual within a decade of active_tol) / all rows


def elem_err(got, ref, floor_frac=1e-3):
    """[B] worst component error of each problem, relative to max(|ref_i|, floor_frac max(||ref||_inf, 1))."""
    got = np.asarray(got, np.float64).reshape(len(ref), -1)
    ref = np.asarray(ref, np.float64).reshape(len(ref), -1)
    floor = floor_frac * np.maximum(np.abs(ref).max(axis=1, keepdims=True), 1.0)
    den = np.maximum(np.abs(ref), floor)
    return (np.abs(got - ref) / den).max(axis=1)


def norm_err(got, ref):
    got = np.asarray(got, np.float64).reshape(len(ref), -1)
    ref = np.asarray(ref, np.float64).reshape(len(ref), -1)
    return np.abs(got - ref).max(axis=1) / np.maximum(1.0, np.abs(ref).max(axis=1))


def row_residuals(params, U, mu=None):
    """Residual of every constraint row in the ABI's bit order (include/lmpc.h LMPC_ACTIVE_WORDS):
    [B][24H] = 16H friction rows, 4H lower-bound rows, 4H upper-bound rows (stance value fz_max)."""
    H = params.horizon
    B = len(U)
    f = np.asarray(U, np.float64).reshape(B, H, 4, 3)
    m = (np.full(B, params.mu) if mu is None else np.asarray(mu, np.float64)).reshape(B, 1, 1)
    fx, fy, fz = f[..., 0], f[..., 1], f[..., 2]
    fr = np.stack([fx + m * fz, m * fz - fx, fy + m * fz, m * fz - fy], axis=-1)  # [B][H][4][4]
    return np.concatenate([fr.reshape(B, 16 * H), fz.reshape(B, 4 * H), (params.fz_max - fz).reshape(B, 4 * H)], axis=1)


def active_mismatch(params, got_bits, ref_bits, ref_U, mu=None):
    """(bad, forgiven, rows): rows whose activity differs between the CUDA path and the oracle.
    A differing row whose ORACLE residual lies within a decade of active_tol ([tol/10, 10 tol]) is a
    tie inside the rounding of the tolerance, not a different active set: it is counted as
    `forgiven` (and reported), every other differing row as `bad`."""
    H = params.horizon
    nbits = 24 * H
    diff = np.asarray(got_bits) ^ np.asarray(ref_bits)
    rows = diff.shape[0] * nbits
    idx = np.nonzero(diff.any(axis=1))[0]
    if len(idx) == 0:
        return 0, 0, rows
    d = np.unpackbits(np.ascontiguousarray(diff[idx]).view(np.uint8), axis=1, bitorder="little")[:, :nbits].astype(bool)
    res = np.abs(row_residuals(params, np.asarray(ref_U)[idx], None if mu is None else np.asarray(mu)[idx]))
    tol = params.active_tol if params.active_tol > 0 else 1e-6
    band = (res >= tol / 10) & (res <= 10 * tol)
    forgiven = int((d & band).sum())
    bad = int((d & ~band).sum())
    return bad, forgiven, rows


def compare(params, got, ref, batch=None, rel_tol=REL_TOL, guard_tol=GUARD_TOL, check_active=True, elem_floor_U=1e-3):
    """Full comparison of one CUDA result dict with one oracle result dict; returns a report dict
    and raises AssertionError (with the numbers) on any violated bound.  The first-step GRFs (the
    north-star quantity) always use the 1e-3 element floor; `elem_floor_U` is the floor for the
    full-horizon U, which tests on ill-conditioned workloads (cond(H) ~ 1e8) set to 1e-2."""
    st = np.asarray(got["status"])
    rep = {"problems": int(len(st)), "status_counts": np.bincount(st, minlength=4).tolist()}
    assert (st <= 1).all(), f"unsolved problems: {rep['status_counts']}"
    keys = [k for k in ("grf", "U") if k in got and k in ref and got[k] is not None]
    for k in keys:
        ee, ne = elem_err(got[k], ref[k], elem_floor_U if k == "U" else 1e-3), norm_err(got[k], ref[k])
        rep[f"{k}_elem_max"], rep[f"{k}_norm_max"] = float(ee.max()), float(ne.max())
        exact = st == 0
        if exact.any():
            rep[f"{k}_norm_max_status0"] = float(ne[exact].max())
        if (~exact).any():  # every interior-point-only answer individually
            rep[f"{k}_elem_max_status1"] = float(ee[~exact].max())
        assert ee.max() <= rel_tol, f"{k}: element-wise error {ee.max():.3e} > {rel_tol} (problem {int(ee.argmax())}, status {int(st[ee.argmax()])})"
        if guard_tol is not None and exact.any():
            q99 = float(np.quantile(ne[exact], 0.99)) if exact.sum() >= 100 else float(np.median(ne[exact]))
            rep[f"{k}_norm_q99_status0"] = q99
            assert q99 <= guard_tol, f"{k}: regression guard, 99 % quantile {q99:.3e} > {guard_tol}"
            assert ne[exact].max() <= GUARD_MAX, f"{k}: regression guard {ne[exact].max():.3e} > {GUARD_MAX} (problem {int(np.nonzero(exact)[0][ne[exact].argmax()])})"
    if check_active:
        mu = batch.get("mu") if batch is not None else None
        bad, forgiven, rows = active_mismatch(params, got["active_set"], ref["active_set"], ref["U"], mu)
        rep["active_bad"], rep["active_forgiven"], rep["active_rows"] = bad, forgiven, rows
        assert bad == 0, f"{bad} constraint rows differ in activity (forgiven ties: {forgiven})"
        assert forgiven <= BAND_FRAC_MAX * rows, f"{forgiven} tie rows forgiven of {rows}"
    return rep
