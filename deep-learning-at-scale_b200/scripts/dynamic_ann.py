"""Dynamic ANN flow regressor (Readme.md:19): the MLP sees a window of --lags consecutive rows
(lag-major, oldest first); the target is the last row's.  Rows are taken in file order as one series.
Usage (cnn.py:2):  python3 dynamic_ann.py columnNames columnTypes targetColumn storagePath [--flags]"""
import _common as C
import dlas
from dlas import ops

MDL_NAME = "dynamic_ann"

if __name__ == "__main__":
    ap = C.parser(__doc__)
    ap.add_argument("--lags", type=int, default=24)
    ap.add_argument("--hidden", default="256,256")
    args = ap.parse_args()
    rank, world = C.init_distributed(args)
    names, cont = C.schema(args)
    X, y = C.load_rows(args, names, cont)
    W = ops.lag_windows(X, args.lags)                       # (n - lags + 1, lags * F)
    yw = y[args.lags - 1:].contiguous()
    train, val, test = C.shard(C.split(W, yw, args.seed, args.standardize), args.batch_size, rank, world)
    dims = [W.shape[1]] + [int(h) for h in args.hidden.split(",")] + [1]
    model = dlas.Sequential(mode=args.mode, seed=args.seed)
    for a, b in zip(dims[:-2], dims[1:-1]):
        model.add(dlas.Dense(a, b, activation="relu"))
    model.add(dlas.Dense(dims[-2], dims[-1]))
    model.compile(loss=args.loss, optimizer=C.optimizer(args))
    C.train_and_report(model, args, MDL_NAME, train, val, test)
