"""Here (CPU): fit erfc(z) ~ 2^-Q(z), Q(z) = z (c1 + z (c2 + ... )) of degree 5, by iteratively re-weighted least squares
(minimax-like) on z in [0, 5]; prints the coefficients pre-scaled for |x| (z = |x| / sqrt 2) as used by gelu_fast() in
csrc/gemm.cu and the fp32 error of the resulting erf-GELU on [-1e4, 1e4]."""
import numpy as np
from scipy.optimize import least_squares
from scipy.special import erf, erfc

DEG = 5
z = np.linspace(1e-6, 5.0, 6001)
target = erfc(z)


def model(c, zz):
    q = np.zeros_like(zz)
    for ck in c[::-1]:
        q = (q + ck) * zz
    return np.exp2(-q)


A = np.vander(z, DEG + 1, increasing=True)[:, 1:]
c, *_ = np.linalg.lstsq(A * target[:, None], (-np.log2(target)) * target, rcond=None)
w, best = np.ones_like(z), None
for _ in range(60):
    c = least_squares(lambda cc: (model(cc, z) - target) * w, c, method="lm").x
    e = np.abs(model(c, z) - target)
    w = w * (1 + 4 * e / e.max())
    w /= w.mean()
    if best is None or e.max() < best[0]:
        best = (e.max(), c.copy())
err, c = best
cp = [ck * (2 ** -0.5) ** (k + 1) for k, ck in enumerate(c)]
print("max |erfc error| %.3e" % err)
print("coefficients for |x| (c1..c5):", ", ".join("%.9ef" % v for v in cp))
x = np.concatenate([np.linspace(-8, 8, 400001), np.linspace(-1e4, 1e4, 200001)]).astype(np.float32)
ax = np.abs(x)
q = np.float32(cp[4])
for ck in cp[3::-1]:
    q = q * ax + np.float32(ck)
pe = np.exp2(-(q * ax)).astype(np.float32)
g = np.float32(0.5) * x * np.where(x < 0, pe, np.float32(2) - pe)
ref = 0.5 * x.astype(np.float64) * (1 + erf(x.astype(np.float64) / np.sqrt(2)))
print("fp32 erf-GELU: max abs error %.3e, all finite %s" % (np.abs(g - ref).max(), bool(np.isfinite(g).all())))
