"""Physical model (Readme.md:7-8): Gilbert's equation q = P*S^1.89 / (435*R^0.546) over the wellhead
records; reports the loss against the measured flow in targetColumn.
Usage (cnn.py:2):  python3 gilbert.py columnNames columnTypes targetColumn storagePath
                   [--pressure-col P --choke-col S --gor-col R] [--data csv | --synthetic N]"""
import time

import _common as C
import torch
import dlas
from dlas import ops

MDL_NAME = "gilbert"

if __name__ == "__main__":
    ap = C.parser(__doc__)
    ap.add_argument("--pressure-col", default="P")
    ap.add_argument("--choke-col", default="S")
    ap.add_argument("--gor-col", default="R")
    args = ap.parse_args()
    t0 = time.time()
    if args.synthetic:
        P, S, R = ops.synth_gilbert(args.synthetic, seed=args.seed)
        q = dlas.gilbert(P, S, R)
        target = None
    else:
        names, cont = C.schema(args)
        X, y = C.load_rows(args, names, cont)
        col = {k: i for i, k in enumerate(cont)}
        P, S, R = (X[:, col[c]].contiguous() for c in (args.pressure_col, args.choke_col, args.gor_col))
        q = dlas.gilbert(P, S, R)
        target = y.reshape(-1)
    torch.cuda.synchronize()
    print("\nTime elapsed: %f s" % (time.time() - t0))                                        # cnn.py:133
    if target is not None:
        d = (q - target).abs()
        loss = d.clamp(max=6.0).mean() if args.loss == "mae_clip" else (d * d).mean()         # cnn.py:29-32
        print("Testing set loss: %f" % float(loss))                                           # cnn.py:134
    else:
        print("rows: %d  mean q: %f stb/d" % (q.numel(), float(q.mean())))
