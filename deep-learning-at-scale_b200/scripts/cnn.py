"""The reference's own model (cnn.py:110-114): Convolution1D(1, 100, 13, relu) -> Dropout(0.5) -> Flatten ->
Dense(3600, 12), mae_clip / SGD, batch 20.  The 3600 = 36 x 100 of the Dense layer fixes the input at 48
samples of one signal; the 12 outputs are read as the next 12 samples: rows are taken in file order, the
target column is the signal, every window of 48 + 12 consecutive values is one (input, target) pair.
Usage (cnn.py:2):  python3 cnn.py columnNames columnTypes targetColumn storagePath [--flags]"""
import _common as C
import dlas
from dlas import ops

MDL_NAME = "cnn"

if __name__ == "__main__":
    ap = C.parser(__doc__)
    ap.add_argument("--filters", type=int, default=100)
    ap.add_argument("--filter-length", type=int, default=13)
    ap.add_argument("--seq-len", type=int, default=48)
    ap.add_argument("--horizon", type=int, default=12)
    ap.add_argument("--dropout", type=float, default=0.5)
    args = ap.parse_args()
    rank, world = C.init_distributed(args)
    names, cont = C.schema(args)
    span = args.seq_len + args.horizon
    if args.synthetic:
        W, _ = ops.synth_series(args.synthetic, span, 1, seed=args.seed)
        W = W.reshape(args.synthetic, span)
    else:
        _, yr = C.load_rows(args, names, cont)
        W = yr.reshape(-1).unfold(0, span, 1)
    X, y = W[:, :args.seq_len].contiguous(), W[:, args.seq_len:].contiguous()
    train, val, test = C.shard(C.split(X, y, args.seed, args.standardize), args.batch_size, rank, world)
    model = dlas.Sequential(seed=args.seed)
    model.add(dlas.Convolution1D(1, args.filters, args.filter_length, activation="relu"))          # cnn.py:111
    model.add(dlas.Dropout(args.dropout))                                                           # cnn.py:112
    model.add(dlas.Flatten())                                                                       # cnn.py:113
    model.add(dlas.Dense((args.seq_len - args.filter_length + 1) * args.filters, args.horizon))     # cnn.py:114
    model.compile(loss=args.loss, optimizer=C.optimizer(args))
    C.train_and_report(model, args, MDL_NAME, train, val, test)
