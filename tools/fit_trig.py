"""Fits the minimax-like polynomials of bx_math.cuh's inline atan2 / sincos / acos (float32) and
reports their max error when evaluated in float32 (Horner with FMA emulated in float64->float32).
usage: python tools/fit_trig.py"""
import numpy as np
from numpy.polynomial import chebyshev as C, polynomial as P


def cheb_fit(f, lo, hi, deg, n=4001, w=None):
  k = np.arange(n)
  x = 0.5 * (lo + hi) + 0.5 * (hi - lo) * np.cos(np.pi * (k + 0.5) / n)
  c = C.Chebyshev.fit(x, f(x), deg, domain=[lo, hi], w=w(x) if w else None)
  return c.convert(kind=P.Polynomial, domain=[-1, 1], window=[-1, 1]).coef


def remez_like(f, lo, hi, deg, iters=30):
  """Lawson-weighted least squares -> near-minimax."""
  n = 8001
  x = np.linspace(lo, hi, n)
  w = np.ones(n)
  for _ in range(iters):
    c = np.polynomial.polynomial.polyfit(x, f(x), deg, w=np.sqrt(w))
    e = np.abs(np.polynomial.polynomial.polyval(x, c) - f(x))
    w = w * (e + 1e-30)
    w /= w.sum()
  return c


def horner32(c, s):
  r = np.full_like(s, np.float32(c[-1]), dtype=np.float32)
  for k in c[-2::-1]:
    r = (r.astype(np.float64) * s.astype(np.float64) + np.float64(np.float32(k))).astype(np.float32)
  return r


def main():
  # atan(a) = a * Pa(a^2), a in [0, 1]
  f = lambda s: np.where(s > 0, np.arctan(np.sqrt(np.maximum(s, 1e-300))) / np.sqrt(np.maximum(s, 1e-300)), 1.0)
  for deg in (6, 7, 8):
    c = remez_like(f, 0.0, 1.0, deg)
    a = np.linspace(0, 1, 200001).astype(np.float32)
    s = (a.astype(np.float64) ** 2).astype(np.float32)
    r = (a.astype(np.float64) * horner32(c, s)).astype(np.float32)
    print('atan deg', deg, 'max err', np.abs(r - np.arctan(a.astype(np.float64))).max(), [float(np.float32(v)) for v in c])
  # sin(x) = x + x^3 * Ps(x^2), cos(x) = 1 + x^2 * Pc(x^2), |x| <= pi/2 (+ margin)
  lim = (np.pi / 2 * 1.02) ** 2
  fs = lambda s: (np.sin(np.sqrt(s)) / np.sqrt(s) - 1) / s
  fc = lambda s: (np.cos(np.sqrt(s)) - 1) / s
  x = np.linspace(-np.pi / 2, np.pi / 2, 200001).astype(np.float32)
  s = (x.astype(np.float64) ** 2).astype(np.float32)
  for deg in (3, 4):
    cs = remez_like(lambda t: fs(np.maximum(t, 1e-12)), 1e-9, lim, deg)
    r = (x.astype(np.float64) + (x.astype(np.float64) * s) * horner32(cs, s)).astype(np.float32)
    print('sin deg', deg, 'max err', np.abs(r - np.sin(x.astype(np.float64))).max(), [float(np.float32(v)) for v in cs])
  for deg in (4, 5):
    cc = remez_like(lambda t: fc(np.maximum(t, 1e-12)), 1e-9, lim, deg)
    r = (1.0 + s.astype(np.float64) * horner32(cc, s)).astype(np.float32)
    print('cos deg', deg, 'max err', np.abs(r - np.cos(x.astype(np.float64))).max(), [float(np.float32(v)) for v in cc])
  # acos(t) = sqrt(1 - t) * Pq(t), t in [0, 1]
  fq = lambda t: np.arccos(np.minimum(t, 1 - 1e-16)) / np.sqrt(1 - np.minimum(t, 1 - 1e-16))
  t = np.linspace(0, 1, 200001).astype(np.float32)
  for deg in (6, 7, 8):
    cq = remez_like(fq, 0.0, 1.0, deg)
    r = (np.sqrt((1.0 - t.astype(np.float64))).astype(np.float32).astype(np.float64) * horner32(cq, t)).astype(np.float32)
    print('acos deg', deg, 'max err', np.abs(r - np.arccos(t.astype(np.float64))).max(), [float(np.float32(v)) for v in cq])


if __name__ == '__main__':
  main()
