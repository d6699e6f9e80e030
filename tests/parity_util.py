"""Prediction parity for the bf16 tensor-core mode: bit-exact, or explained row by row.

The north star asks for bit-exact argmax. In bf16 mode that holds against `oracle/mlp_numpy.py(emulate_bf16=True)` as
long as both sides hold the same weights. After an optimizer step the two weight sets agree only within the 1e-2 band
(bf16 rounding of dH / dZ amplifies fp32 summation-order differences), so a row whose top-2 logits are closer than
that drift can legitimately flip. `assert_predictions_explained` therefore checks TWO statements and never a rate:

  (A) exact under own weights: the engine's predictions equal argmax of the oracle's logits computed from the
      ENGINE's current weights (`own_params`, read back before the step). Allowed exceptions are ties only: rows whose
      top-2 margin under those weights is below `tie_bound` = 1e-5 * max|z_row| (fp32 summation order) + one bf16 ulp
      of the largest activation-weight product (2^-7 * max_k |a_k| * max|W_last|: a hidden activation that lands on
      the other side of a bf16 rounding boundary).
  (B) every row that differs from the ORACLE TRAJECTORY (oracle weights) has a trajectory margin below
      2 * max_c |z_own - z_traj|[row] + tie_bound, i.e. the measured logit drift between the two weight sets explains
      it; and that drift itself stays within the weight tolerance: max|z_own - z_traj| <= 1e-2 * max|z_traj|.

Rows excepted under (A) or (B) are returned (and printed with their margins); every other row is identical, so correct
counts match exactly once those rows are excluded."""
from __future__ import annotations

import numpy as np

from oracle import mlp_numpy as onp

TIE_REL = 1e-5
DRIFT_REL = 1e-2


def _margins(z: np.ndarray, rows: np.ndarray) -> np.ndarray:
    top2 = np.sort(z[rows], axis=1)[:, -2:]
    return top2[:, 1] - top2[:, 0]


def _tie_bound(z: np.ndarray, acts_last: np.ndarray, w_last: np.ndarray, rows: np.ndarray) -> np.ndarray:
    return TIE_REL * np.abs(z[rows]).max(axis=1) + 2.0 ** -7 * np.abs(acts_last[rows]).max(axis=1) * np.abs(w_last).max()


def assert_predictions_explained(got, oracle: onp.OracleMLP, own_params, x, masks, tag: str, y=None, n_correct=None):
    """`oracle` must not have taken the step yet (its weights are the trajectory's pre-step weights)."""
    got = np.asarray(got)
    z_traj, acts_t, _ = oracle._forward(x, masks)
    own = onp.OracleMLP(weights=own_params[0::2], biases=own_params[1::2], dropout=oracle.dropout, activation=oracle.activation,
                        emulate_bf16=oracle.emulate_bf16)
    z_own, acts_o, _ = own._forward(x, masks)
    # (A) exact under the engine's own weights, ties excepted
    bad_a = np.nonzero(got != z_own.argmax(axis=1))[0]
    if len(bad_a):
        m = _margins(z_own, bad_a)
        tb = _tie_bound(z_own, acts_o[-1], own.weights[-1], bad_a)
        print(f'[parity] {tag}: {len(bad_a)} rows differ under the engine\'s own weights; margins {m[:8]} (tie bounds {tb[:8]})')
        assert (m <= tb).all(), f'{tag}: predictions differ from argmax under the engine\'s own weights beyond a tie: rows {bad_a[:8]}, ' \
                                f'margins {m[:8]}, bounds {tb[:8]}'
        top2 = np.argsort(z_own[bad_a], axis=1)[:, -2:]
        assert all(g in t for g, t in zip(got[bad_a], top2)), f'{tag}: a flipped row is not a top-2 class'
    # (B) differences from the oracle trajectory are explained by the measured logit drift
    drift_row = np.abs(z_own - z_traj).max(axis=1)
    scale = float(np.abs(z_traj).max())
    assert drift_row.max() <= DRIFT_REL * scale, f'{tag}: logit drift {drift_row.max():.3e} beyond {DRIFT_REL} * {scale:.3f}'
    bad_b = np.nonzero(got != z_traj.argmax(axis=1))[0]
    if len(bad_b):
        m = _margins(z_traj, bad_b)
        bound = 2.0 * drift_row[bad_b] + _tie_bound(z_traj, acts_t[-1], oracle.weights[-1], bad_b)
        print(f'[parity] {tag}: {len(bad_b)} of {len(got)} rows differ from the oracle trajectory; margins {m[:8]}, '
              f'bounds {bound[:8]} (max drift {drift_row.max():.3e})')
        assert (m <= bound).all(), f'{tag}: trajectory mismatches not explained by weight drift: rows {bad_b[:8]}, margins {m[:8]}, ' \
                                   f'bounds {bound[:8]}'
    excepted = np.union1d(bad_a, bad_b)
    if y is not None and n_correct is not None:
        # the fused count is exactly the count of the engine's own predictions; outside the excepted rows it is the oracle's
        y = np.asarray(y)
        assert int(n_correct) == int((got == y).sum()), f'{tag}: fused correct count {n_correct} != count of its own predictions'
        keep = np.ones(len(got), dtype=bool)
        keep[excepted] = False
        assert int((got[keep] == y[keep]).sum()) == int((z_traj.argmax(axis=1)[keep] == y[keep]).sum())
    return excepted


def bf16_noise_margin(oracle_f32: onp.OracleMLP, x, masks) -> np.ndarray:
    """Per-row 6-sigma bound on the logit change caused by rounding inputs, weights and hidden activations to bf16 (each
    rounding an independent relative error, uniform within half an ulp = 2^-9: variance (2^-9)^2 / 3), for one hidden
    layer nets. Used when a bf16 run is compared with an fp32 run of the reference: rows whose fp32 top-2 margin is
    below this can flip."""
    assert len(oracle_f32.weights) == 2
    v = (2.0 ** -9) ** 2 / 3.0
    xf = np.asarray(x, dtype=np.float64).reshape(x.shape[0], -1)
    w1, w2 = oracle_f32.weights[0].astype(np.float64), oracle_f32.weights[1].astype(np.float64)
    _, acts, _ = oracle_f32._forward(x, masks)
    a = acts[-1].astype(np.float64)
    var_a = 2.0 * v * ((xf * xf) @ (w1 * w1).T) + v * a * a           # x and W1 roundings, then the activation's own
    var_z = var_a @ (w2 * w2).T + 2.0 * v * ((a * a) @ (w2 * w2).T)   # ... and W2's
    return 6.0 * np.sqrt(2.0 * var_z.max(axis=1))


class LockstepF32:
    """Walks the pinned fp32 oracle next to a bf16 engine through a golden replay (tests/golden_util.replay interface:
    numpy in, numpy out). A prediction may differ from the fp32 trajectory only where the fp32 top-2 margin is below that
    row's bf16 rounding noise (6 sigma) plus twice the logit drift between the fp32 trajectory's weights and the engine's
    fp32 master weights (`own_params()`), both evaluated in fp32. `flips` collects (margin, bound) of the rows that did."""

    def __init__(self, engine, own_params, f32: onp.OracleMLP):
        self.engine, self.own_params, self.f32, self.flips = engine, own_params, f32, []

    def _bound(self, x, masks):
        z, _, _ = self.f32._forward(x, masks)
        own = self.own_params()
        z_own, _, _ = onp.OracleMLP(weights=own[0::2], biases=own[1::2], dropout=self.f32.dropout,
                                    activation=self.f32.activation)._forward(x, masks)
        return z, bf16_noise_margin(self.f32, x, masks) + 2.0 * np.abs(z_own - z).max(axis=1)

    def _check(self, p, z, bound):
        bad = np.nonzero(np.asarray(p) != z.argmax(axis=1))[0]
        if len(bad):
            m = _margins(z, bad)
            self.flips.extend(zip(m.tolist(), bound[bad].tolist()))
            assert (m <= bound[bad]).all(), f'bf16 vs fp32 reference: margins {m} above the bf16 noise bounds {bound[bad]}'

    def fit(self, x, y, masks):
        z, bound = self._bound(x, masks)
        p, l = self.engine.fit(x, y, masks)
        self.f32.fit(x, y, masks)
        self._check(p, z, bound)
        return p, l

    def evaluate(self, x, y):
        z, bound = self._bound(x, None)
        p, l = self.engine.evaluate(x, y)
        self._check(p, z, bound)
        return p, l


def relu_gate_ties(oracle: onp.OracleMLP, x, masks, rel: float = 2e-6) -> set:
    """Hidden units of the FIRST hidden layer whose pre-activation sits at fp32 round-off distance from zero for some
    row of this batch (|u| < rel * max|u|): a different summation order can put such a value on the other side of ReLU's
    gate, which switches that row's whole contribution to the unit's weight row on or off. Same nature as an argmax tie;
    returned so the weight comparison can name the units it treats loosely (tests report them)."""
    _, _, pre = oracle._forward(x, masks)
    u = pre[0]
    live = np.ones_like(u, dtype=bool) if masks is None or oracle.dropout == 0.0 else masks[0].astype(bool)
    near = (np.abs(u) < rel * np.abs(u).max()) & live
    return set(np.nonzero(near.any(axis=0))[0].tolist())
