"""Device generators, split codes and lag windows against the oracle (bit-exact: integer hash and
exactly-rounded fp32 operations only)."""
import numpy as np
import pytest
import torch

import dlas_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    import __graft_entry__ as G
    G.build()
    from dlas import ops
    return ops


def test_synth_gilbert_bit_exact(ops):
    P, S, R = ops.synth_gilbert(100_003, start=999)
    Ph, Sh, Rh = O.synth_gilbert(O.BASE_SEED, 999, 100_003)
    assert np.array_equal(P.cpu().numpy(), Ph) and np.array_equal(S.cpu().numpy(), Sh) and np.array_equal(R.cpu().numpy(), Rh)


@pytest.mark.parametrize("F", [8, 32])
def test_synth_rows_bit_exact(ops, F):
    X, y = ops.synth_rows(20_001, F, start=77)
    Xh, yh = O.synth_rows(O.BASE_SEED, 77, 20_001, F)
    assert np.array_equal(X.cpu().numpy(), Xh)
    yd = y.cpu().numpy().ravel()
    assert np.mean(yd == yh) > 0.9999 and np.abs(yd - yh).max() < 2e-7  # emulated-fma double rounding only


def test_synth_series_bit_exact(ops):
    X, y = ops.synth_series(513, 48, 8, start=5)
    Xh, yh = O.synth_series(O.BASE_SEED, 5, 513, 48, 8)
    assert np.mean(X.cpu().numpy() == Xh) > 0.9999 and np.abs(X.cpu().numpy() - Xh).max() < 2e-7
    assert np.abs(y.cpu().numpy().ravel() - yh).max() < 1e-6


def test_split_codes(ops):
    c = ops.split_codes(300_000, start=11).cpu().numpy()
    assert np.array_equal(c, O.split_code(O.BASE_SEED, np.arange(11, 300_011)))


def test_lag_windows(ops):
    s, _ = O.synth_rows(O.BASE_SEED, 0, 1000, 8)
    w = ops.lag_windows(torch.from_numpy(s).cuda(), 24).cpu().numpy()
    assert np.array_equal(w, O.lag_windows(s, 24))
    assert ops.lag_windows(torch.from_numpy(s[:5]).cuda(), 24).shape == (0, 192)
