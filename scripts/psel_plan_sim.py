import numpy as np, sys
from scipy.stats import norm
rng = np.random.default_rng(2)

def gen(n, npx, hot=1e-3, cold=2e-4):
    x = rng.standard_normal((npx, n))
    m = rng.random((npx, n)) < hot; x[m] += rng.uniform(20, 60, m.sum())
    m = rng.random((npx, n)) < cold; x[m] -= 5
    return x

def clip_final(x, s=3.0):
    keep = np.ones_like(x, bool)
    for it in range(30):
        cnt = keep.sum(1)
        xm = np.where(keep, x, np.nan)
        sd = np.nanstd(xm, axis=1, ddof=1); med = np.nanmedian(xm, axis=1)
        nk = keep & ~((med[:,None]-x > sd[:,None]*s) | (x-med[:,None] > sd[:,None]*s))
        if (nk == keep).all(): break
        keep = nk
    return keep, med, sd

def sim(n, npx, h, WRc, mfk, trim):
    x = gen(n, npx)
    S1 = x.sum(1); S2 = (x*x).sum(1); nn = np.full(npx, n, float)
    if trim:
        mx = x.max(1); mn = x.min(1)
        S1 = S1 - mx - mn; S2 = S2 - mx*mx - mn*mn; nn = nn - 2
    mean = S1/nn; sd = np.sqrt((S2 - S1*S1/nn)/(nn-1))
    hw = h*sd
    f = (np.abs(x-mean[:,None]) < hw[:,None]).mean(1)
    sat = f >= 0.95
    hw2 = np.where(sat, hw/4, hw)
    f = (np.abs(x-mean[:,None]) < hw2[:,None]).mean(1)
    f = np.clip(f, 0.05, 0.95)
    srob = hw2/norm.ppf((1+f)/2)
    q = np.clip(((x < mean[:,None]).sum(1)+0.5)/n, 0.03, 0.97)
    p1 = mean - srob*norm.ppf(q)
    WR = max(6, WRc*n**0.25)
    w = WR*srob*2.5066/n
    mf = 1/(1+4*mfk/np.sqrt(n))
    r = mf*3*srob
    keep, med, sdf = clip_final(x)
    a = (~keep & (x < med[:,None])).sum(1); b = (~keep & (x > med[:,None])).sum(1)
    cl = (x <= (p1-w)[:,None]).sum(1); cc = (np.abs(x-p1[:,None]) < w[:,None]).sum(1)
    klo = (n-1)//2 - (b+1)//2 - 1; khi = n//2 + (a+1)//2 + 1
    miss_c = ~((cl <= klo) & (cl+cc-1 >= khi))
    miss_t = ((p1-r) < (med-3*sdf)) | ((p1+r) > (med+3*sdf))
    nt = (np.abs(x-p1[:,None]) > r[:,None]).sum(1)
    return dict(n=n,h=h,WRc=WRc,mfk=mfk,trim=trim, miss_c=miss_c.mean(), miss_t=miss_t.mean(), cc=cc.mean(), ccmax=cc.max(), nt=nt.mean(), ntmax=nt.max(), sat=sat.mean(), srob_sd=np.std(srob), srob_mean=np.mean(srob))

n = int(sys.argv[1]); npx = int(sys.argv[2])
for h in (0.35, 0.6745, 1.0):
  for trim in (0, 1):
    for mfk in (1.7, 2.5):
      print(sim(n, npx, h, 2.5, mfk, trim))
