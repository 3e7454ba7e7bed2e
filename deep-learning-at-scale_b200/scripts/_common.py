"""Shared driver of the spark-submit-style scripts.  Same positional contract as the reference
(cnn.py:2, :41-44):  script.py columnNames columnTypes targetColumn storagePath  [optional --flags]
columnNames / columnTypes are comma-separated; types follow cnn.py:53-59 ("int", "float", anything
else = string).  String columns are not on the B200 hot path (their one-hot ETL is Spark-side,
cnn.py:72-91) and are ignored with a notice.  The target column is EXCLUDED from the features
(the reference leaks it).  Rows come from --data <headerless csv> (cnn.py:65) or
--synthetic N (counter-hash generator); the 64/16/20 split (cnn.py:68) is a seeded hash of the row."""
from __future__ import annotations

import argparse
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import torch  # noqa: E402
import dlas  # noqa: E402
from dlas import ops  # noqa: E402


def parser(description):
    ap = argparse.ArgumentParser(description=description)
    ap.add_argument("columnNames")
    ap.add_argument("columnTypes")
    ap.add_argument("targetColumn")
    ap.add_argument("storagePath")
    ap.add_argument("--data", help="header-less CSV with the declared columns (cnn.py:65)")
    ap.add_argument("--synthetic", type=int, default=0, help="generate N synthetic rows instead of reading --data")
    ap.add_argument("--epochs", type=int, default=1000)          # cnn.py:128 nb_epoch=1000
    ap.add_argument("--batch-size", type=int, default=20)        # cnn.py:128 batch_size=20
    ap.add_argument("--patience", type=int, default=10)          # cnn.py:121
    ap.add_argument("--loss", default="mae_clip", choices=["mae_clip", "mse"])   # cnn.py:118
    ap.add_argument("--optimizer", default="sgd", choices=["sgd", "adam"])
    ap.add_argument("--mode", default="fp32", choices=["fp32", "tc"])
    ap.add_argument("--seed", type=int, default=ops.BASE_SEED)
    return ap


def schema(args):
    names, types = args.columnNames.split(","), args.columnTypes.split(",")
    if len(names) != len(types):
        raise SystemExit("columnNames and columnTypes differ in length")
    print([(k, v if v in ("int", "float") else "string") for k, v in zip(names, types)])   # cnn.py:62 prints the schema
    cat = [k for k, v in zip(names, types) if v not in ("int", "float")]
    print("Categorical variables: " + ", ".join(cat))                                      # cnn.py:73
    if cat:
        print("notice: string columns are ignored on the GPU path (one-hot encoding is Spark-side ETL)")
    cont = [k for k, v in zip(names, types) if v in ("int", "float") and k != args.targetColumn]
    if args.targetColumn not in names:
        raise SystemExit("targetColumn %r is not among columnNames" % args.targetColumn)
    return names, cont


def load_rows(args, names, cont):
    """Returns device tensors X (n, len(cont)) and y (n, 1)."""
    if args.synthetic:
        return ops.synth_rows(args.synthetic, len(cont), seed=args.seed)
    if not args.data:
        raise SystemExit("give --data <csv> or --synthetic N")
    import pandas as pd
    df = pd.read_csv(args.data, header=None, names=names)
    X = torch.as_tensor(df[cont].to_numpy(dtype=np.float32)).cuda()
    y = torch.as_tensor(df[[args.targetColumn]].to_numpy(dtype=np.float32)).cuda()
    return X, y


def split(X, y, seed):
    code = ops.split_codes(X.shape[0], seed=seed)
    parts = []
    for k in range(3):
        idx = (code == k).nonzero(as_tuple=True)[0]
        parts.append((X.index_select(0, idx).contiguous(), y.index_select(0, idx).contiguous()))
    return parts


def optimizer(args):
    if args.optimizer == "adam":
        return dlas.Adam()
    return dlas.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True)   # cnn.py:117


def train_and_report(model, args, name, train, val, test):
    callbacks = [dlas.EarlyStopping(monitor="val_loss", patience=args.patience),                       # cnn.py:121
                 dlas.ModelCheckpoint(args.storagePath + "models/%s.mdl" % name, save_best_only=True)]  # cnn.py:122
    t0 = time.time()
    model.fit(train[0], train[1], validation_data=val, verbose=2, callbacks=callbacks, nb_epoch=args.epochs,
              batch_size=args.batch_size)                                                               # cnn.py:127-128
    torch.cuda.synchronize()
    time_elapsed = time.time() - t0
    test_set_loss = model.test_on_batch(test[0], test[1])                                               # cnn.py:132
    print("\nTime elapsed: %f s" % time_elapsed)                                                        # cnn.py:133
    print("Testing set loss: %f" % test_set_loss)                                                       # cnn.py:134
    return time_elapsed, test_set_loss
