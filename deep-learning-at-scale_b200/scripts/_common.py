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
    ap.add_argument("--gpus", type=int, default=1,
                    help="data-parallel training on G GPUs of this box: the script re-launches itself as G ranks "
                         "(torch.distributed.run); under torchrun the ranks are taken from the environment")
    ap.add_argument("--master-port", type=int, default=29541)
    ap.add_argument("--standardize", action="store_true",
                    help="z-score every feature column with the training split's mean / deviation before training (the "
                         "reference feeds raw columns, cnn.py:53-65; --mode tc keeps rows as fp16 operands and needs "
                         "columns within +-65504, see DLAS_MODE_TC in include/dlas_b200.h)")
    ap.add_argument("--one-hot", action="store_true",
                    help="with --data: encode string columns as cnn.py:72-91 does (StringIndexer -> OneHotEncoder) and put "
                         "them in front of the continuous columns (cnn.py:93-99); host-side, fitted once on all rows")
    return ap


def init_distributed(args):
    """One process per GPU.  `--gpus G` outside torchrun re-launches this script as G ranks on 127.0.0.1;
    inside a rank: bind the GPU, join the NCCL process group, and leave printing to rank 0.  Returns (rank, world)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world == 1 and getattr(args, "gpus", 1) > 1:
        import subprocess
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", str(args.master_port), sys.argv[0]] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    if world == 1:
        return 0, 1
    import torch.distributed as dist
    rank, local = int(os.environ["RANK"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    from dlas import parallel
    parallel.bind_host_to_gpu(local)           # host buffers first-touched on the GPU's own NUMA node
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if rank != 0:
        sys.stdout = open(os.devnull, "w")      # the reference prints once (cnn.py:62,:73,:133-134): rank 0 does
    return rank, world


def shard(parts, batch_size, rank, world):
    """This rank's rows of the train / validation / test parts.  Train: the rank's contiguous 1/world slice of every global
    batch (dlas.parallel.local_rows), and the same share of the ragged tail (at most world-1 rows are dropped so that every
    rank runs the same steps); validation / test: every world-th row (evaluate() pools sums and counts over the ranks)."""
    if world == 1:
        return parts
    from dlas import parallel
    (Xt, yt), val, test = parts
    n = Xt.shape[0]
    n_whole = n // batch_size * batch_size
    idx = parallel.local_rows(n_whole, batch_size, rank, world).to(Xt.device)
    tail = (n - n_whole) // world
    if tail:
        idx = torch.cat([idx, torch.arange(n_whole + rank * tail, n_whole + (rank + 1) * tail, device=Xt.device)])
    out = [(Xt.index_select(0, idx).contiguous(), yt.index_select(0, idx).contiguous())]
    for X, y in (val, test):
        out.append((X[rank::world].contiguous(), y[rank::world].contiguous()))
    return out


def schema(args):
    names, types = args.columnNames.split(","), args.columnTypes.split(",")
    if len(names) != len(types):
        raise SystemExit("columnNames and columnTypes differ in length")
    print([(k, v if v in ("int", "float") else "string") for k, v in zip(names, types)])   # cnn.py:62 prints the schema
    cat = [k for k, v in zip(names, types) if v not in ("int", "float")]
    print("Categorical variables: " + ", ".join(cat))                                      # cnn.py:73
    if cat and not getattr(args, "one_hot", False):
        print("notice: string columns are ignored (pass --one-hot to encode them as cnn.py:72-91 does)")
    cont = [k for k, v in zip(names, types) if v in ("int", "float") and k != args.targetColumn]
    if args.targetColumn not in names:
        raise SystemExit("targetColumn %r is not among columnNames" % args.targetColumn)
    return names, cont


def string_index(values):
    """Spark StringIndexer (cnn.py:75), default order: most frequent label first, ties by label ascending."""
    labels, counts = np.unique(np.asarray(values, dtype=str), return_counts=True)
    order = sorted(range(len(labels)), key=lambda i: (-counts[i], labels[i]))
    rank = {labels[i]: r for r, i in enumerate(order)}
    return np.fromiter((rank[str(v)] for v in values), dtype=np.int64, count=len(values)), [labels[i] for i in order]


def one_hot_drop_last(index, n_labels):
    """Spark OneHotEncoder (cnn.py:77-80), default dropLast=True: n_labels - 1 columns, the last label is all zeros."""
    out = np.zeros((len(index), max(0, n_labels - 1)), dtype=np.float32)
    keep = index < n_labels - 1
    out[np.nonzero(keep)[0], index[keep]] = 1.0
    return out


def assemble_features(df, names, types, target, one_hot=False):
    """VectorAssembler order of cnn.py:93-99: the encoded categorical columns first (argv order), then the continuous
    columns in argv order -- WITHOUT the target (the reference leaks it).  The reference fits the
    indexers separately on the train / validation / test splits (cnn.py:88-90), which gives the three splits different
    encodings; here they are fitted once on all rows.  Returns (X float32 (n, F), y float32 (n, 1))."""
    cont = [k for k, v in zip(names, types) if v in ("int", "float") and k != target]
    cat = [k for k, v in zip(names, types) if v not in ("int", "float") and k != target]
    blocks = []
    if one_hot:
        for c in cat:
            idx, labels = string_index(df[c].to_numpy())
            blocks.append(one_hot_drop_last(idx, len(labels)))
    blocks.append(df[cont].to_numpy(dtype=np.float32).reshape(len(df), len(cont)))
    X = np.ascontiguousarray(np.concatenate(blocks, axis=1), dtype=np.float32)
    return X, df[[target]].to_numpy(dtype=np.float32)


def load_rows(args, names, cont):
    """Returns device tensors X (n, len(cont)) and y (n, 1)."""
    if args.synthetic:
        return ops.synth_rows(args.synthetic, len(cont), seed=args.seed)
    if not args.data:
        raise SystemExit("give --data <csv> or --synthetic N")
    import pandas as pd
    df = pd.read_csv(args.data, header=None, names=names)
    Xh, yh = assemble_features(df, names, args.columnTypes.split(","), args.targetColumn, one_hot=getattr(args, "one_hot", False))
    return torch.as_tensor(Xh).cuda(), torch.as_tensor(yh).cuda()


def split(X, y, seed, standardize=False):
    code = ops.split_codes(X.shape[0], seed=seed)
    parts = []
    for k in range(3):
        idx = (code == k).nonzero(as_tuple=True)[0]
        parts.append((X.index_select(0, idx).contiguous(), y.index_select(0, idx).contiguous()))
    if standardize and parts[0][0].shape[0] > 1:     # --standardize: statistics of the training split only (host-side ETL)
        mu = parts[0][0].mean(dim=0, keepdim=True)
        sd = parts[0][0].std(dim=0, keepdim=True).clamp_min(1e-12)
        parts = [(((Xp - mu) / sd).contiguous(), yp) for Xp, yp in parts]
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
    if getattr(model, "world", 1) > 1 or getattr(getattr(model, "_impl", None), "world", 1) > 1:
        # mapped peer memory and NCCL communicators: leave without the collective teardown (as bench.py does)
        import torch.distributed as dist
        sys.stdout.flush()
        torch.cuda.synchronize()
        dist.barrier()
        os._exit(0)
    return time_elapsed, test_set_loss
