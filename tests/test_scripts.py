"""The spark-submit-style scripts keep the reference's argv contract (cnn.py:2, :41-44, :53-59) and its
two closing print lines (cnn.py:133-134)."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCRIPTS = os.path.join(ROOT, "deep-learning-at-scale_b200", "scripts")


def test_schema_rule_and_target_exclusion(capsys):
    sys.path.insert(0, SCRIPTS)
    import _common as C
    ap = C.parser("t")
    args = ap.parse_args(["a,b,well,q", "float,int,string,float", "q", "/tmp/store/"])
    names, cont = C.schema(args)
    out = capsys.readouterr().out
    assert names == ["a", "b", "well", "q"] and cont == ["a", "b"]          # target excluded (fixes D5), strings ignored
    assert "Categorical variables: well" in out                             # cnn.py:73
    assert (args.epochs, args.batch_size, args.patience, args.loss) == (1000, 20, 10, "mae_clip")   # cnn.py:121,128,118
    with pytest.raises(SystemExit):
        C.schema(ap.parse_args(["a,b", "float", "a", "/tmp/"]))


@pytest.mark.gpu
@pytest.mark.parametrize("script,extra", [
    ("static_ann.py", ["--synthetic", "20000", "--hidden", "64,64", "--batch-size", "256", "--epochs", "3", "--loss", "mse"]),
    ("dynamic_ann.py", ["--synthetic", "6000", "--lags", "24", "--hidden", "256,256", "--batch-size", "256", "--epochs", "2"]),
    ("lstm.py", ["--synthetic", "600", "--seq-len", "12", "--hidden", "32", "--layers", "2", "--batch-size", "64", "--epochs", "2"]),
    ("cnn.py", ["--synthetic", "800", "--batch-size", "20", "--epochs", "2"]),
    ("gilbert.py", ["--synthetic", "100000"]),
])
def test_scripts_run_with_reference_argv(tmp_path, script, extra):
    names = ",".join("f%d" % i for i in range(8)) + ",q"
    types = ",".join(["float"] * 9)
    store = str(tmp_path) + "/"
    r = subprocess.run([sys.executable, os.path.join(SCRIPTS, script), names, types, "q", store] + extra,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    assert "Time elapsed:" in r.stdout
    if script != "gilbert.py":
        assert "Testing set loss:" in r.stdout and "Epoch 1/" in r.stdout
        assert os.path.exists(os.path.join(store, "models", script[:-3] + ".mdl"))          # cnn.py:122


@pytest.mark.gpu
@pytest.mark.parametrize("one_hot", [False, True])
def test_static_ann_from_csv(tmp_path, one_hot):
    rng = np.random.default_rng(0)
    n = 5000
    A = rng.normal(size=(n, 3)).astype(np.float32)
    q = (A @ np.float32([0.5, -1.0, 2.0]) + 0.1).astype(np.float32)
    path = str(tmp_path / "wells.csv")
    with open(path, "w") as f:
        for i in range(n):
            f.write("%g,%g,w%d,%g,%g\n" % (A[i, 0], A[i, 1], i % 7, A[i, 2], q[i]))
    r = subprocess.run([sys.executable, os.path.join(SCRIPTS, "static_ann.py"), "a,b,well,c,q", "float,float,string,float,float", "q",
                        str(tmp_path) + "/", "--data", path, "--hidden", "64,64", "--batch-size", "128", "--epochs", "30",
                        "--optimizer", "adam", "--loss", "mse"] + (["--one-hot"] if one_hot else []), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    loss = float(r.stdout.split("Testing set loss:")[1].split()[0])
    assert loss < 0.05 * float(np.var(q))      # it learns the linear map


def test_feature_assembly_follows_the_reference_pipeline():
    """cnn.py:72-99 restated on the host: StringIndexer (most frequent label first), OneHotEncoder (last label dropped),
    categorical block in front of the continuous columns, target excluded (fixes D5)."""
    import pandas as pd
    sys.path.insert(0, SCRIPTS)
    import _common as C
    df = pd.DataFrame({"a": [1.0, 2.0, 3.0, 4.0, 5.0, 6.0], "well": ["w2", "w1", "w2", "w3", "w2", "w1"],
                       "n": [1, 2, 3, 4, 5, 6], "q": [10.0, 20.0, 30.0, 40.0, 50.0, 60.0]})
    names, types = ["a", "well", "n", "q"], ["float", "string", "int", "float"]
    idx, labels = C.string_index(df["well"].to_numpy())
    assert labels == ["w2", "w1", "w3"] and idx.tolist() == [0, 1, 0, 2, 0, 1]          # by frequency, descending
    assert C.string_index(np.array(["b", "a", "a", "b"]))[1] == ["a", "b"]             # ties: label ascending
    oh = C.one_hot_drop_last(idx, 3)
    assert oh.shape == (6, 2) and oh.tolist() == [[1, 0], [0, 1], [1, 0], [0, 0], [1, 0], [0, 1]]
    X, y = C.assemble_features(df, names, types, "q", one_hot=True)
    assert X.dtype == np.float32 and X.shape == (6, 4) and y.shape == (6, 1)
    np.testing.assert_array_equal(X[:, :2], oh)                                          # categorical block first (cnn.py:96)
    np.testing.assert_array_equal(X[:, 2], df["a"]); np.testing.assert_array_equal(X[:, 3], df["n"])
    np.testing.assert_array_equal(y[:, 0], df["q"])                                      # the target is not a feature
    X0, _ = C.assemble_features(df, names, types, "q", one_hot=False)
    np.testing.assert_array_equal(X0, X[:, 2:])                                          # default: numeric columns only
    Xt, yt = C.assemble_features(df, names, types, "a", one_hot=True)                    # any numeric column can be the target
    assert Xt.shape == (6, 4) and yt[:, 0].tolist() == df["a"].tolist()


def test_shard_gives_every_rank_the_same_steps():
    """scripts/_common.shard (--gpus G): rank r owns the r-th contiguous 1/G slice of every global batch and the same share
    of the ragged tail, so every rank runs the same number of steps of the same sizes (a precondition of model.fit)."""
    import torch
    sys.path.insert(0, SCRIPTS)
    import _common as C
    n, B, W = 1000, 96, 4                     # 10 whole batches of 96 + a tail of 40 rows
    X = torch.arange(n, dtype=torch.float32).reshape(n, 1)
    parts = [(X, X.clone()), (X[:50], X[:50]), (X[:31], X[:31])]
    got = [C.shard(parts, B, r, W) for r in range(W)]
    assert {g[0][0].shape[0] for g in got} == {10 * (B // W) + 40 // W}
    for b in range(10):                        # the union of the ranks' slices of batch b is batch b, in rank order
        rows = torch.cat([g[0][0][b * 24:(b + 1) * 24, 0] for g in got])
        assert rows.tolist() == list(range(b * B, (b + 1) * B))
    tail = torch.cat([g[0][0][240:, 0] for g in got])
    assert tail.tolist() == list(range(960, 1000))
    assert sum(g[1][0].shape[0] for g in got) == 50 and sum(g[2][0].shape[0] for g in got) == 31
    assert C.shard(parts, B, 0, 1) is parts


def _wellhead_csv(path, n, big_column=False):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import dlas_oracle as O
    P, S, R = O.synth_gilbert(O.BASE_SEED, 0, n)
    cols = [P, S, R]
    X = np.stack([cols[j % 3] * (1.0 + 0.01 * (j // 3)) for j in range(32)], axis=1)
    if big_column:
        X[:, 5] = 1.0e5 + 20.0 * X[:, 5]
    q = np.log10(O.gilbert(P, S, R))
    np.savetxt(path, np.concatenate([X, q[:, None]], axis=1), delimiter=",", fmt="%.7g")
    names = ",".join("c%d" % i for i in range(32)) + ",q"
    return names, ",".join(["float"] * 33)


@pytest.mark.gpu
def test_static_ann_tc_mode_on_wellhead_scale_columns(tmp_path):
    """Raw wellhead columns (P 50..5000 psig, S 4..128, R 0.02..10) through the drop-in script in the tensor-core mode:
    with --standardize the fp16-operand run ends where the fp32 run does; a column beyond the fp16 range without it is
    refused loudly (DLAS_ERR_RANGE), and the fp32 mode takes the same file."""
    path = str(tmp_path / "wells.csv")
    names, types = _wellhead_csv(path, 30000)
    common = ["--data", path, "--hidden", "256,256", "--batch-size", "1024", "--epochs", "4", "--optimizer", "adam", "--loss", "mse"]
    losses = {}
    for mode in ("fp32", "tc"):
        r = subprocess.run([sys.executable, os.path.join(SCRIPTS, "static_ann.py"), names, types, "q", str(tmp_path) + "/" + mode + "/"]
                           + common + ["--mode", mode, "--standardize"], capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
        losses[mode] = float(r.stdout.split("Testing set loss:")[1].split()[0])
    assert np.isfinite(losses["tc"]) and losses["tc"] == pytest.approx(losses["fp32"], rel=2e-2), losses
    names, types = _wellhead_csv(path, 8000, big_column=True)
    r = subprocess.run([sys.executable, os.path.join(SCRIPTS, "static_ann.py"), names, types, "q", str(tmp_path) + "/big/"] + common
                       + ["--mode", "tc", "--epochs", "1"], capture_output=True, text=True, timeout=600)
    assert r.returncode != 0 and "65504" in (r.stdout + r.stderr), r.stdout[-800:] + r.stderr[-800:]
    r = subprocess.run([sys.executable, os.path.join(SCRIPTS, "static_ann.py"), names, types, "q", str(tmp_path) + "/big32/"] + common
                       + ["--mode", "fp32", "--epochs", "1", "--standardize"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]


@pytest.mark.gpu
def test_dynamic_ann_tc_mode_matches_fp32(tmp_path):
    """The dynamic-ANN script (24 lags x 8 features = 192-256-256-1, Readme.md:17-19) in the tensor-core mode: layer 1 in
    mlp_tc_l1.cu around the CTA-pair kernel (the script materialises the windows: its random row split breaks the series); the run
    ends where the fp32 run does."""
    names = ",".join("f%d" % i for i in range(8)) + ",q"
    types = ",".join(["float"] * 9)
    losses = {}
    for mode in ("fp32", "tc"):
        r = subprocess.run([sys.executable, os.path.join(SCRIPTS, "dynamic_ann.py"), names, types, "q", str(tmp_path) + "/" + mode + "/",
                            "--synthetic", "40000", "--lags", "24", "--hidden", "256,256", "--batch-size", "2048", "--epochs", "4",
                            "--loss", "mse", "--mode", mode], capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
        losses[mode] = float(r.stdout.split("Testing set loss:")[1].split()[0])
    assert np.isfinite(losses["tc"]) and losses["tc"] == pytest.approx(losses["fp32"], rel=1e-2), losses
