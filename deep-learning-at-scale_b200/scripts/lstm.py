"""LSTM flow regressor (Readme.md:21): --layers stacked LSTM(--hidden) over windows of --seq-len rows,
last hidden state -> Dense(1).  Rows are taken in file order as one series.
Usage (cnn.py:2):  python3 lstm.py columnNames columnTypes targetColumn storagePath [--flags]"""
import _common as C
import dlas
from dlas import ops

MDL_NAME = "lstm"

if __name__ == "__main__":
    ap = C.parser(__doc__)
    ap.add_argument("--seq-len", type=int, default=48)
    ap.add_argument("--hidden", type=int, default=128)
    ap.add_argument("--layers", type=int, default=2)
    args = ap.parse_args()
    rank, world = C.init_distributed(args)
    names, cont = C.schema(args)
    if args.synthetic:
        X, y = ops.synth_series(args.synthetic, args.seq_len, len(cont), seed=args.seed)
        X = X.reshape(args.synthetic, -1)
    else:
        Xr, yr = C.load_rows(args, names, cont)
        X = ops.lag_windows(Xr, args.seq_len)               # each window is one (T, F) sequence, batch-major
        y = yr[args.seq_len - 1:].contiguous()
    train, val, test = C.shard(C.split(X, y, args.seed, args.standardize), args.batch_size, rank, world)
    model = dlas.Sequential(seed=args.seed)
    for i in range(args.layers):
        model.add(dlas.LSTM(args.hidden, input_shape=(args.seq_len, len(cont)) if i == 0 else None,
                            return_sequences=(i < args.layers - 1)))
    model.add(dlas.Dense(1))
    model.compile(loss=args.loss, optimizer=C.optimizer(args))
    C.train_and_report(model, args, MDL_NAME, train, val, test)
