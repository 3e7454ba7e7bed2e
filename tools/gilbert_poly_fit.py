"""Design-time script: fit the log2 / exp2 polynomials used by csrc/gilbert.cu (near-minimax by
Lawson-reweighted least squares on Chebyshev nodes, fp64) and print them rounded to fp32.
No GPU needed.  Run: python tools/gilbert_poly_fit.py"""
import numpy as np

def cheb_nodes(a, b, n):
    k = np.arange(n)
    return 0.5*(a+b) + 0.5*(b-a)*np.cos(np.pi*(2*k+1)/(2*n))

def fit(fun, a, b, deg, iters=60, npts=3000):
    x = cheb_nodes(a, b, npts); y = fun(x)
    t = (2*x-(a+b))/(b-a)                       # well-conditioned basis
    V = np.polynomial.chebyshev.chebvander(t, deg)
    w = np.ones_like(x)
    for _ in range(iters):
        c, *_ = np.linalg.lstsq(V*np.sqrt(w)[:,None], y*np.sqrt(w), rcond=None)
        err = np.abs(V@c - y)
        w = w*(err/err.max() + 1e-2); w /= w.sum()
    # convert to monomial in x
    p = np.polynomial.chebyshev.Chebyshev(c, domain=[a,b]).convert(kind=np.polynomial.Polynomial)
    return p.coef, err.max()

def q_fun(f):   # (log2(1+f)/f - 1/ln2)/f by its series: (1/ln2) * sum_{k>=2} (-1)^(k+1) f^(k-2)/k
    s = np.zeros_like(f)
    for k in range(200, 1, -1):
        s = s*f + (-1.0)**(k+1)/k
    return s/np.log(2.0)

lo, hi = np.sqrt(0.5)-1, np.sqrt(2.0)-1
if __name__ == "__main__":
    for deg in (6,7,8,9):
        c, e = fit(q_fun, lo, hi, deg)
        print("log2 Q deg", deg, "abs err of log2 <=", e*hi*hi)
        if deg in (7,8): print("  coef fp32:", [float(np.float32(v)) for v in c])
    for deg in (5,6):
        c, e = fit(lambda r: np.exp2(r), -0.5, 0.5, deg)
        print("exp2 deg", deg, "abs err", e, "\n  coef fp32:", [float(np.float32(v)) for v in c])


# ---------------------------------------------------------------------------------------------
# fp32 emulation of the kernel algorithm (mirrors csrc/gilbert.cu, poly variant)
# ---------------------------------------------------------------------------------------------
F = np.float32
def fma(a, b, c):
    return (np.asarray(a, np.float64)*np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(np.float32)

C0 = F(1.4426950408889634)
def q_fun32(f):   # refit against the fp32-rounded leading coefficient
    return q_fun(f) + (1/np.log(2.0) - float(C0))/np.where(np.abs(f) < 1e-12, 1e-12, f) * 0  # (difference folded below)

def split(x):
    i = x.view(np.int32)
    e = (i - np.int32(0x3f3504f3)) >> 23
    m = (i - (e << 23)).view(np.float32)
    return e.astype(np.float32), (m - F(1.0)).astype(np.float32)

def log2m(f, QC):
    q = np.full_like(f, QC[-1])
    for c in QC[-2::-1]:
        q = fma(q, f, c)
    u = fma(f, q, C0)
    return (f*u).astype(np.float32)

def exp2r(r, EC):
    p = np.full_like(r, EC[-1])
    for c in EC[-2::-1]:
        p = fma(p, r, c)
    return p

C1, C2 = F(1.89), F(0.546)
C1H, C2H, C0H = F(121/64), F(35/64), F(561/64)
C1L, C2L, C0L = F(1.89 - 121/64), F(0.546 - 35/64), F(np.log2(435.0) - 561/64)

def gilbert32(P, S, R, QC, EC):
    eP, fP = split(P); eS, fS = split(S); eR, fR = split(R)
    LP, LS, LR = log2m(fP, QC), log2m(fS, QC), log2m(fR, QC)
    th = fma(-C2H, eR, fma(C1H, eS, (eP - C0H).astype(np.float32)))      # exact multiples of 1/64
    tl = fma(-C2L, eR, fma(C1L, eS, -C0L))
    tf = fma(-C2, LR, fma(C1, LS, LP))
    ts = (tl + tf).astype(np.float32)
    n = np.rint((th + ts).astype(np.float32))
    r = ((th - n).astype(np.float32) + ts).astype(np.float32)
    p = exp2r(r, EC)
    return np.ldexp(p, n.astype(np.int32)).astype(np.float32)

if __name__ == "__main__":
    QC, _ = fit(q_fun, lo, hi, 7); QC = QC.astype(np.float32)
    EC, _ = fit(lambda r: np.exp2(r), -0.5, 0.5, 6); EC = EC.astype(np.float32)
    rng = np.random.default_rng(1)
    worst = 0.0
    for rep in range(8):
        n = 4_000_000
        P = (50 + 4950*rng.random(n)).astype(np.float32)
        S = (4 + 124*rng.random(n)).astype(np.float32)
        R = (0.02*500**rng.random(n)).astype(np.float32)
        if rep >= 6:   # wide dynamic range
            P = np.exp(rng.uniform(-20, 20, n)).astype(np.float32)
            S = np.exp(rng.uniform(-10, 10, n)).astype(np.float32)
            R = np.exp(rng.uniform(-10, 10, n)).astype(np.float32)
        q = gilbert32(P, S, R, QC, EC)
        ref = P.astype(np.float64)*S.astype(np.float64)**1.89/(435.0*R.astype(np.float64)**0.546)
        rel = np.abs(q/ref - 1).max()
        worst = max(worst, rel); print(rep, "max rel err", rel)
    print("worst", worst)
    print("QC =", [float(v) for v in QC]); print("EC =", [float(v) for v in EC])
