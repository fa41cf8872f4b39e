"""Fits q(u) with 2^q(u) ~ Phi(-u) = erfc(u / sqrt 2) / 2 on [0, 5.5], minimising max |u (2^q(u) - Phi(-u))| -- the absolute
error of gelu(x) = relu(x) - |x| 2^q(|x|) (csrc/common.cuh: gelu_erf).  Prints degree, max error, coefficients (highest first)."""
import numpy as np
from scipy.special import erfc
from scipy.optimize import least_squares
U=5.5
u=np.linspace(0,U,20001)
E=0.5*erfc(u/np.sqrt(2))
L=np.log2(E)
def run(deg):
    # initial: LS fit of log2 E
    c=np.polyfit(u,L,deg)
    w=np.ones_like(u)
    def err(c): return u*(np.exp2(np.polyval(c,u))-E)
    # Lawson-like reweighting on abs error
    best=None
    for it in range(60):
        r=least_squares(lambda c: w*err(c), c, method='lm', xtol=1e-15, ftol=1e-15)
        c=r.x
        e=np.abs(err(c))
        m=e.max()
        if best is None or m<best[0]: best=(m,c.copy())
        w=w*(0.5+e/m); w/=w.mean()
    return best
for deg in (4,5,6,7):
    m,c=run(deg)
    print(deg, m, list(c))
