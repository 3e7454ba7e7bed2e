"""GPU parity of the Gilbert kernel (through the C ABI) against the fp64 oracle.  Tolerance from
BASELINE.json: <= 1e-6 relative in fp32, measured against fp64 on the fp32-rounded inputs."""
import os

import numpy as np
import pytest
import torch

import dlas_oracle as O

pytestmark = pytest.mark.gpu
TOL = 1e-6
VARIANTS = [0, 7, 5, 2]  # poly; all-MUFU; MUFU for P,R only; MUFU for S only


@pytest.fixture(scope="module")
def ops():
    import __graft_entry__ as G
    G.build()
    from dlas import ops
    return ops


def _run(ops, P, S, R, variant):
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    q = ops.gilbert(dev(P), dev(S), dev(R), variant=variant)
    torch.cuda.synchronize()
    return q.cpu().numpy()


@pytest.mark.parametrize("variant", VARIANTS)
def test_golden_kat(ops, golden_dir, variant):
    g = np.load(os.path.join(golden_dir, "gilbert_kat.npz"))
    q = _run(ops, g["P"], g["S"], g["R"], variant)
    ref = g["q64"]
    nz = ref != 0
    assert np.abs(q[nz] / ref[nz] - 1).max() <= TOL
    assert np.all(q[~nz] == 0)                       # P == 0 -> q == 0
    assert q[0] == pytest.approx(1.0, rel=TOL)       # P=435, S=1, R=1


@pytest.mark.parametrize("variant", VARIANTS)
def test_bench_distribution_4m(ops, variant):
    P, S, R = O.synth_gilbert(O.BASE_SEED, 10_000_000, 1 << 22)
    q = _run(ops, P, S, R, variant)
    rel = np.abs(q / O.gilbert(P, S, R) - 1)
    print("variant %d max rel %.3g mean %.3g" % (variant, rel.max(), rel.mean()))
    assert rel.max() <= TOL


def test_ragged_unaligned_and_empty(ops):
    P, S, R = O.synth_gilbert(O.BASE_SEED, 0, 1003)
    ref = O.gilbert(P, S, R)
    for n in (0, 1, 3, 5, 1003):
        q = _run(ops, P[:n], S[:n], R[:n], 0)
        assert q.shape == (n,)
        if n:
            assert np.abs(q / ref[:n] - 1).max() <= TOL
    # unaligned views take the scalar kernel
    Pd, Sd, Rd = (torch.from_numpy(a).cuda() for a in (P, S, R))
    q = ops.gilbert(Pd[1:].clone()[0:], Sd[1:].contiguous(), Rd[1:].contiguous())
    out = torch.empty(1003, device="cuda")[1:]
    ops.gilbert(Pd[1:], Sd[1:], Rd[1:], out=out)
    torch.cuda.synchronize()
    assert np.abs(out.cpu().numpy() / ref[1:] - 1).max() <= TOL


def test_special_inputs_take_the_slow_path(ops):
    f = np.float32
    P = f([0.0, -1.0, np.nan, np.inf, 1e-40, 100.0, 100.0, 3e38, 1e-30])
    S = f([16.0, 16.0, 16.0, 16.0, 16.0, 0.0, np.inf, 1e30, 1e-20])
    R = f([1.0, 1.0, 1.0, 1.0, 1.0, 1.0, 1.0, 1e-30, 1e30])
    q = _run(ops, P, S, R, 0)
    with np.errstate(all="ignore"):
        ref = (P.astype(np.float64) * S.astype(np.float64) ** 1.89 / (435.0 * R.astype(np.float64) ** 0.546)).astype(f)
    # P == 0 -> 0; negative P is just a negative flow; NaN propagates; inf stays inf
    assert q[0] == 0 and q[1] == pytest.approx(ref[1], rel=1e-5) and np.isnan(q[2]) and np.isinf(q[3])
    assert q[5] == 0 and np.isinf(q[6]) and np.isinf(q[7]) and q[8] == 0
    assert q[4] == pytest.approx(ref[4], rel=1e-4)    # subnormal input: powf path
    # negative choke size or gas-oil ratio has no real power: NaN, as the fp64 formula gives
    q = _run(ops, f([100.0, 100.0]), f([-4.0, 16.0]), f([1.0, -2.0]), 0)
    assert np.isnan(q).all()


def test_full_size_properties(ops):
    """200M rows (BASELINE.json config 2): homogeneity q(2P) = 2 q(P) to 2 ulp (a power-of-two scale
    only moves the exponent; the split point of 2^t may move with it), S-monotonicity on a sorted
    choke sweep, plus oracle spot slices."""
    n = 200_000_000
    P, S, R = ops.synth_gilbert(n)
    q = ops.gilbert(P, S, R)
    q2 = ops.gilbert(P * 2, S, R)
    assert float(((q2 - q * 2).abs() / (q * 2)).max()) <= 2.4e-7
    for start in (0, 123_456_789, n - 4096):
        Ph, Sh, Rh = O.synth_gilbert(O.BASE_SEED, start, 4096)
        assert np.array_equal(P[start:start + 4096].cpu().numpy(), Ph)
        rel = np.abs(q[start:start + 4096].cpu().numpy() / O.gilbert(Ph, Sh, Rh) - 1).max()
        assert rel <= TOL
    del q2
    Ss, _ = torch.sort(S[: 1 << 24])
    qs = ops.gilbert(torch.full_like(Ss, 1000.0), Ss, torch.full_like(Ss, 0.7))
    assert bool((qs[1:] >= qs[:-1]).all())
