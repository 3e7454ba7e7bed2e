"""Static ANN flow regressor (Readme.md:17): dense MLP on per-well tabular features.
Usage (cnn.py:2):  python3 static_ann.py columnNames columnTypes targetColumn storagePath [--flags]"""
import _common as C
import dlas

MDL_NAME = "static_ann"

if __name__ == "__main__":
    ap = C.parser(__doc__)
    ap.add_argument("--hidden", default="256,256", help="hidden layer widths (BASELINE.json: 64,64 or 256,256)")
    args = ap.parse_args()
    rank, world = C.init_distributed(args)
    names, cont = C.schema(args)
    X, y = C.load_rows(args, names, cont)
    train, val, test = C.shard(C.split(X, y, args.seed, args.standardize), args.batch_size, rank, world)
    dims = [int(X.shape[1])] + [int(h) for h in args.hidden.split(",")] + [1]   # len(cont), plus the one-hot columns with --one-hot
    model = dlas.Sequential(mode=args.mode, seed=args.seed)
    for a, b in zip(dims[:-2], dims[1:-1]):
        model.add(dlas.Dense(a, b, activation="relu"))
    model.add(dlas.Dense(dims[-2], dims[-1]))
    model.compile(loss=args.loss, optimizer=C.optimizer(args))
    C.train_and_report(model, args, MDL_NAME, train, val, test)
