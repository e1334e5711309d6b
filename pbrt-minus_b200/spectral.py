"""Host-side spectral data for the scene front-end (the reference builds the same objects at start-up):

  GlobalSpectra::create          spectrum_util/global_spectra.cu:7-52   CIE XYZ, normalised D65
  RGBColorSpace                  spectrum_util/rgb_color_space.h:11-42  sRGB primaries, whitepoint from D65
  RGBtoSpectrumTable             spectrum_util/rgb_to_spectrum_data.cu  Jakob-Hanika sigmoid fit, 64^3 x 3 table
  PiecewiseLinearSpectrum(...)   spectra/piecewise_linear_spectrum.cu:5-35 (interleaved + normalise)

The reference fits the WHOLE 3 x 64^3 coefficient table on the CPU at every start-up ("spectra
computing took ... seconds") and then interpolates it trilinearly per RGB value.  Here only the
table cells a lookup touches are fitted (same Gauss-Newton iteration, same warm-start chain along
the brightness axis, same trilinear interpolation), so loading a scene costs milliseconds.
"""
import functools
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LAMBDA_MIN, LAMBDA_MAX = 360, 830
CIE_Y_INTEGRAL = np.float32(106.856895)
RES = 64


@functools.lru_cache(maxsize=1)
def tables():
    return dict(np.load(os.path.join(HERE, "data", "tables.npz")))


def _f32(a):
    return np.asarray(a, dtype=np.float32)


def piecewise_eval(lam, val, x):
    """PiecewiseLinearSpectrum::operator() at float wavelengths x (0 outside the knots)."""
    lam = np.asarray(lam, dtype=np.float64)
    val = np.asarray(val, dtype=np.float64)
    x = np.asarray(x, dtype=np.float64)
    y = np.interp(x, lam, val)
    return np.where((x < lam[0]) | (x > lam[-1]), 0.0, y)


def piecewise_eval32(lam, val, x):
    """PiecewiseLinearSpectrum::operator() (spectra/piecewise_linear_spectrum.cu:93-103) in the reference's fp32 arithmetic, one
    rounding per operation: t = (x - l0) / (l1 - l0), then lerp = (1 - t) * a + t * b (util/math.h:74-76).  The colour space's
    illuminant and every light's photometric scale come out of this, so it has to round the way the reference's host does."""
    lam, val = _f32(lam), _f32(val)
    x = _f32(x)
    n = len(lam)
    o = np.clip(np.searchsorted(lam, x, side="right") - 1, 0, n - 2)     # find_interval: the last knot <= x, clamped to [0, n - 2]
    l0, l1, a, b = lam[o], lam[o + 1], val[o], val[o + 1]
    t = ((x - l0).astype(np.float32) / (l1 - l0).astype(np.float32)).astype(np.float32)
    lo = ((np.float32(1) - t).astype(np.float32) * a).astype(np.float32)
    hi = (t * b).astype(np.float32)
    y = (lo + hi).astype(np.float32)
    return np.where((x < lam[0]) | (x > lam[-1]), np.float32(0), y).astype(np.float32)


def _fma32(a, b, c):
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(np.float32)


def sigmoid_poly32(c, lam):
    """RGBSigmoidPolynomial::operator() (spectrum_util/rgb_sigmoid_polynomial.h:14-37) as the reference's HOST rounds it:
    Horner with fma, then .5 + x / (2 sqrt(1 + x^2)) one fp32 operation at a time"""
    lam = _f32(lam)
    c0, c1, c2 = (np.float32(v) for v in c[:3])
    x = _fma32(lam, _fma32(lam, c0, c1), c2)
    with np.errstate(invalid="ignore", over="ignore"):
        xx = (x * x).astype(np.float32)
        root = np.sqrt((np.float32(1) + xx).astype(np.float32)).astype(np.float32)
        y = (np.float32(0.5) + (x / (np.float32(2) * root).astype(np.float32)).astype(np.float32)).astype(np.float32)
    return np.where(np.isinf(x), np.where(x > 0, np.float32(1), np.float32(0)), y).astype(np.float32)


def spectrum_eval32(rec, lam):
    """Spectrum::operator()(lambda) of a front-end spectrum record in the reference's host fp32 arithmetic (types as in
    include/b200pt.h: constant, rgb albedo / unbounded / illuminant, piecewise linear, dense)"""
    lam = _f32(lam)
    t, c = rec["type"], rec["c"]
    if t == 0:
        return np.full(lam.shape, np.float32(c[0]), dtype=np.float32)
    if t == 1:
        return sigmoid_poly32(c, lam)
    if t == 2:
        return (np.float32(c[3]) * sigmoid_poly32(c, lam)).astype(np.float32)
    if t == 3:
        a = (np.float32(c[3]) * sigmoid_poly32(c, lam)).astype(np.float32)
        grid = np.arange(LAMBDA_MIN, LAMBDA_MAX + 1).astype(np.float32)
        return (a * piecewise_eval32(grid, illuminant_d65()[2], lam)).astype(np.float32)
    if t == 4:
        return piecewise_eval32(rec["lambda"], rec["value"], lam)
    if t == 5:
        grid = np.arange(LAMBDA_MIN, LAMBDA_MAX + 1).astype(np.float32)
        return piecewise_eval32(grid, rec["value"], lam)
    raise ValueError("spectrum type %r" % t)


def dense_scaled32(rec, scale):
    """DenselySampledSpectrum(spectrum, scale) (spectra/densely_sampled_spectrum.h:10-14): spectrum(lambda) * scale at 360..830"""
    grid = np.arange(LAMBDA_MIN, LAMBDA_MAX + 1).astype(np.float32)
    return (spectrum_eval32(rec, grid) * np.float32(scale)).astype(np.float32)


def inner_product32(a, b):
    """Spectrum::inner_product (e.g. spectra/densely_sampled_spectrum.h:17-24): an fp32 running sum of fp32 products over the
    integer wavelengths 360..830"""
    acc = np.float32(0)
    for p, q in zip(_f32(a), _f32(b)):
        acc = np.float32(acc + np.float32(p * q))
    return acc


def interleaved_to_knots(samples):
    """spectra/piecewise_linear_spectrum.cu:5-27: extend to [360, 830] with constant ends."""
    s = np.asarray(samples, dtype=np.float32)
    lam, val = list(s[0::2]), list(s[1::2])
    if lam[0] > LAMBDA_MIN:
        lam.insert(0, np.float32(LAMBDA_MIN - 1))
        val.insert(0, val[0])
    if lam[-1] < LAMBDA_MAX:
        lam.append(np.float32(LAMBDA_MAX + 1))
        val.append(val[-1])
    return _f32(lam), _f32(val)


@functools.lru_cache(maxsize=1)
def cie_xyz():
    t = tables()
    return _f32(t["cie_x"]), _f32(t["cie_y"]), _f32(t["cie_z"])


@functools.lru_cache(maxsize=1)
def illuminant_d65():
    """The colour-space illuminant: CIE D65 (5 nm knots), scaled so that its inner product with CIE Y
    is CIE_Y_integral; returned as knots and as the dense 471-sample table the device uses (a
    piecewise-linear spectrum with knots on integer nm is reproduced exactly by 1-nm lerp)."""
    lam, val = interleaved_to_knots(tables()["illum_d65_interleaved"])
    grid = np.arange(LAMBDA_MIN, LAMBDA_MAX + 1)
    y = cie_xyz()[1]
    acc = inner_product32(piecewise_eval32(lam, val, grid), y)
    scale = np.float32(np.float32(CIE_Y_INTEGRAL) / acc)
    val = _f32(val * scale)
    dense = piecewise_eval32(lam, val, grid)
    return lam, val, dense


def inner_product_dense(a, b):
    return float(np.sum(_f32(a).astype(np.float64) * _f32(b).astype(np.float64)))


def spectrum_to_xyz(dense):
    x, y, z = cie_xyz()
    return np.array([inner_product_dense(dense, x), inner_product_dense(dense, y), inner_product_dense(dense, z)]) / float(CIE_Y_INTEGRAL)


@functools.lru_cache(maxsize=1)
def srgb_color_space():
    """xyz_from_rgb / rgb_from_xyz of RGBColorSpace (sRGB primaries, whitepoint = the illuminant's xy)."""
    W = spectrum_to_xyz(illuminant_d65()[2])
    wx, wy = W[0] / W.sum(), W[1] / W.sum()

    def from_xy(x, y, Y=1.0):
        return np.array([x * Y / y, Y, (1 - x - y) * Y / y])

    r, g, b = from_xy(0.64, 0.33), from_xy(0.3, 0.6), from_xy(0.15, 0.06)
    rgb = np.stack([r, g, b], axis=1)
    Wp = from_xy(wx, wy)
    c = np.linalg.solve(rgb, Wp)
    xyz_from_rgb = rgb @ np.diag(c)
    return {"xyz_from_rgb": xyz_from_rgb, "rgb_from_xyz": np.linalg.inv(xyz_from_rgb), "w": (wx, wy)}


# ---------------------------------------------------------------------------------------------
# RGB -> sigmoid-polynomial coefficients (spectrum_util/rgb_to_spectrum_data.cu)
_XYZ_TO_SRGB = np.array([[3.240479, -1.537150, -0.498535], [-0.969256, 1.875991, 0.041556], [0.055648, -0.204043, 1.057311]])
_SRGB_TO_XYZ = np.array([[0.412453, 0.357580, 0.180423], [0.212671, 0.715160, 0.072169], [0.019334, 0.119193, 0.950227]])
_FINE = (95 - 1) * 3 + 1


@functools.lru_cache(maxsize=1)
def _fit_tables():
    """init_tables (:172-212): Simpson-3/8 weights folded with CIE XYZ * D65 -> rgb_tbl, whitepoint."""
    t = tables()
    h = (LAMBDA_MAX - LAMBDA_MIN) / (_FINE - 1)
    lam = LAMBDA_MIN + np.arange(_FINE) * h

    def interp(data, x):
        x = (x - LAMBDA_MIN) * (95 - 1) / (LAMBDA_MAX - LAMBDA_MIN)
        off = np.clip(x.astype(np.int64), 0, 95 - 2)
        w = x - off
        return (1.0 - w) * data[off] + w * data[off + 1]

    xyz = np.stack([interp(t["fit_cie_x"], lam), interp(t["fit_cie_y"], lam), interp(t["fit_cie_z"], lam)])
    I = interp(t["fit_cie_d65"], lam)
    i = np.arange(_FINE)
    ratio = np.where((i == 0) | (i == _FINE - 1), 1.0, np.where((i - 1) % 3 == 2, 2.0, 3.0))
    weight = 3.0 / 8.0 * h * ratio
    rgb_tbl = _XYZ_TO_SRGB @ (xyz * I * weight)
    white = (xyz * I * weight).sum(axis=1)
    lam01 = (lam - LAMBDA_MIN) / (LAMBDA_MAX - LAMBDA_MIN)
    return rgb_tbl, white, lam01


def _cie_lab(p, white):
    X, Y, Z = _SRGB_TO_XYZ @ p

    def f(t):
        d = 6.0 / 29.0
        return np.cbrt(t) if t > d * d * d else t / (d * d * 3.0) + 4.0 / 29.0

    fy = f(Y / white[1])
    return np.array([116.0 * fy - 16.0, 500.0 * (f(X / white[0]) - fy), 200.0 * (fy - f(Z / white[2]))])


def _residual(coeffs, rgb):
    rgb_tbl, white, lam01 = _fit_tables()
    x = (coeffs[0] * lam01 + coeffs[1]) * lam01 + coeffs[2]
    s = 0.5 * x / np.sqrt(1.0 + x * x) + 0.5
    out = rgb_tbl @ s
    return _cie_lab(np.asarray(rgb, dtype=np.float64), white) - _cie_lab(out, white)


def _gauss_newton(rgb, coeffs, iters=15):
    eps = 1e-4
    coeffs = np.array(coeffs, dtype=np.float64)
    for _ in range(iters):
        r = _residual(coeffs, rgb)
        J = np.zeros((3, 3))
        for i in range(3):
            tmp = coeffs.copy()
            tmp[i] -= eps
            r0 = _residual(tmp, rgb)
            tmp = coeffs.copy()
            tmp[i] += eps
            r1 = _residual(tmp, rgb)
            J[:, i] = (r1 - r0) / (2 * eps)
        coeffs = coeffs - np.linalg.solve(J, r)
        mx = coeffs.max()
        if mx > 200:
            coeffs *= 200 / mx
        if float(r @ r) < 1e-6:
            break
    return coeffs


def _smoothstep(x):
    return x * x * (3.0 - 2.0 * x)


_Z_NODES = np.array([_smoothstep(_smoothstep(k / (RES - 1))) for k in range(RES)])


@functools.lru_cache(maxsize=4096)
def _column(l, j, i):
    """compute() (:305-359) for one (l, j, i) column: all RES brightness levels, warm-started outwards
    from k = RES/5 exactly as the reference chains them.  Returns [RES, 3] polynomial coefficients in nm."""
    y, x = j / (RES - 1), i / (RES - 1)
    c0, c1 = 360.0, 1.0 / (830.0 - 360.0)
    start = RES // 5
    out = np.zeros((RES, 3))

    def store(k, c):
        A, B, C = c
        out[k] = [A * c1 * c1, B * c1 - 2 * A * c0 * c1 * c1, C - B * c0 * c1 + A * (c0 * c1) ** 2]

    for ks in (range(start, RES), range(start, -1, -1)):
        coeffs = np.zeros(3)
        for k in ks:
            b = _Z_NODES[k]
            rgb = np.zeros(3)
            rgb[l] = b
            rgb[(l + 1) % 3] = x * b
            rgb[(l + 2) % 3] = y * b
            coeffs = _gauss_newton(rgb, coeffs)
            store(k, coeffs)
    return out


def rgb_to_sigmoid(rgb):
    """RGBtoSpectrumTable::operator() (:387-427): trilinear lookup in the (lazily fitted) table."""
    rgb = np.clip(np.asarray(rgb, dtype=np.float32), 0.0, np.inf)
    r, g, b = (float(v) for v in rgb)
    if r == g == b:
        v = np.float32(r)
        return _f32([0.0, 0.0, (v - np.float32(0.5)) / np.sqrt(np.float64(v) * (1.0 - np.float64(v)))])
    m = (0 if r > b else 2) if r > g else (1 if g > b else 2)
    c = [r, g, b]
    z = c[m]
    # the reference's mixed arithmetic: rgb * (RES - 1) is an fp32 product, x / y / z and the offsets are doubles, and every lerp
    # is pbrt::lerp(Real, Real, Real) -- its double arguments (offsets, table entries) are rounded to fp32 at the call and the
    # interpolation itself runs in fp32 on the host (no contraction)
    x = float(np.float32(np.float32(c[(m + 1) % 3]) * np.float32(RES - 1))) / z
    y = float(np.float32(np.float32(c[(m + 2) % 3]) * np.float32(RES - 1))) / z
    xi, yi = min(int(x), RES - 2), min(int(y), RES - 2)
    zi = int(np.clip(np.searchsorted(_Z_NODES, z, side="left") - 1, 0, RES - 2))
    dx, dy = x - xi, y - yi
    dz = (z - _Z_NODES[zi]) / (_Z_NODES[zi + 1] - _Z_NODES[zi])

    def lerp(t, a, b_):
        t, a, b_ = np.float32(t), np.asarray(a, np.float32), np.asarray(b_, np.float32)
        lo = ((np.float32(1) - t).astype(np.float32) * a).astype(np.float32)
        return (lo + (t * b_).astype(np.float32)).astype(np.float32)

    def co(ddx, ddy, ddz):
        return _column(m, yi + ddy, xi + ddx)[zi + ddz]

    out = lerp(dz, lerp(dy, lerp(dx, co(0, 0, 0), co(1, 0, 0)), lerp(dx, co(0, 1, 0), co(1, 1, 0))),
               lerp(dy, lerp(dx, co(0, 0, 1), co(1, 0, 1)), lerp(dx, co(0, 1, 1), co(1, 1, 1))))
    return _f32(out)


NAMED = {"metal-Ag-eta": "Ag_eta", "metal-Ag-k": "Ag_k", "metal-Al-eta": "Al_eta", "metal-Al-k": "Al_k",
         "metal-Au-eta": "Au_eta", "metal-Au-k": "Au_k", "metal-Cu-eta": "Cu_eta", "metal-Cu-k": "Cu_k",
         "glass-BK7": "GlassBK7_eta", "glass-F11": "GlassSF11_eta"}


def named_spectrum(name):
    """scene/scene_builder.cu:100-136: named metal / glass spectra as piecewise-linear knots."""
    if name not in NAMED:
        raise KeyError("unknown named spectrum '%s'" % name)
    return interleaved_to_knots(tables()["named_" + NAMED[name]])
