"""oracle/f2f_oracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

numpy restatement of the reference's frame-to-frame step
(blind_denoising.py:44-122, twin lib/frame2frame/warped_loss.py:117-248) with
the arithmetic of the third-party ops it calls restated as executed by the
installed torch 2.11 CPU build (probed in this container):

* `F.grid_sample(bilinear, zeros, align_corners=False)` -- ATen's vectorised CPU
  kernel unnormalises with one fused multiply-add, ix = fma(g+1, W/2, -0.5),
  takes weights e = 1-w, s = 1-n, nw = s*e ..., and blends with a chain of
  fused multiply-adds fma(se_v,se, fma(sw_v,sw, fma(ne_v,ne, nw_v*nw)));
* `scipy.ndimage.binary_dilation` with the 3x3 cross, one iteration,
  border_value 0.

Pinned against the reference itself (oracle/f2f_ref.py, run in the build
container) by tests/test_oracle_f2f.py and tests/golden/f2f_step_ref.npz.
Only tests/, smoke() and bench.py's cpu_baseline leg may import this.
"""
import numpy as np

f32 = np.float32


def _fma(a, b, c):
    # fused multiply-add on float32 arrays: the float64 product of two float32 is exact
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(np.float32)


def warp(x, flo, align_corners=False):
    """blind_denoising.py:49-78.  x [H,W] float32, flo [2,H,W] float32 ->
    (warped [H,W] float32, valid [H,W] {0,1} float32)."""
    H, W = x.shape
    yy, xx = np.mgrid[0:H, 0:W]
    X = xx.astype(f32) + flo[0]                       # :62-65 grid.float() + flo
    Y = yy.astype(f32) + flo[1]
    gx = (f32(2.0) * X) / f32(max(W - 1, 1)) - f32(1.0)   # :68-69
    gy = (f32(2.0) * Y) / f32(max(H - 1, 1)) - f32(1.0)
    if align_corners:
        ix = (gx + f32(1)) * f32((W - 1) / 2)
        iy = (gy + f32(1)) * f32((H - 1) / 2)
    else:                                            # the as-run default
        ix = _fma(gx + f32(1), f32(W / 2), f32(-0.5))
        iy = _fma(gy + f32(1), f32(H / 2), f32(-0.5))
    x0, y0 = np.floor(ix), np.floor(iy)
    w = ix - x0
    e = f32(1) - w
    n = iy - y0
    s = f32(1) - n
    nw, ne, sw, se = s * e, s * w, n * e, n * w
    xi, yi = x0.astype(np.int64), y0.astype(np.int64)

    def tap(img, yq, xq):
        m = (xq >= 0) & (xq < W) & (yq >= 0) & (yq < H)
        v = np.zeros((H, W), f32)
        v[m] = img[yq[m], xq[m]]
        return v

    def blend(img):
        a, b, c, d = tap(img, yi, xi), tap(img, yi, xi + 1), tap(img, yi + 1, xi), tap(img, yi + 1, xi + 1)
        return _fma(d, se, _fma(c, sw, _fma(b, ne, a * nw)))

    warped = blend(x.astype(f32))
    wsum = blend(np.ones((H, W), f32))
    valid = (wsum >= f32(0.9999)).astype(f32)         # :75-76
    return warped, valid


def occlusion(flo, f64=True):
    """blind_denoising.py:88-109 (f64=True) / warped_loss.py:184-207 (f64=False):
    occluded-pixel map {0,1} uint8 including the cross dilation and the 1-px frame."""
    H, W = flo.shape[1:]
    if f64:
        a = np.zeros((H, W), np.float64)
        b = np.zeros((H, W), np.float64)
    else:
        a = np.zeros((H, W), f32)
        b = np.zeros((H, W), f32)
    a[:-1, :] = flo[0, 1:, :] - flo[0, :-1, :]
    b[:, :-1] = flo[1, :, 1:] - flo[1, :, :-1]
    occ = np.abs(a + b) > (0.75 if f64 else f32(0.75))
    d = occ.copy()
    d[1:, :] |= occ[:-1, :]
    d[:-1, :] |= occ[1:, :]
    d[:, 1:] |= occ[:, :-1]
    d[:, :-1] |= occ[:, 1:]
    d[0, :] = d[-1, :] = True
    d[:, 0] = d[:, -1] = True
    return d.astype(np.uint8)


def warp_and_mask(prev, flo, f64=True, align_corners=False):
    warped, valid = warp(prev, flo, align_corners)
    mask = valid * (f32(1) - occlusion(flo, f64).astype(f32))      # :112
    return warped, valid, mask


def masked_loss(inp, warped, mask, kind="l1sum"):
    """blind_denoising.py:121 (kind='l1sum': nn.L1Loss(size_average=False)), and the lib
    variants warped_loss.py:227,239-248 ('l2' mean of squares, 'l1' mean sqrt(d^2+1e-6)).
    Returns (loss float64, dL/d inp float32)."""
    d = (mask * inp - mask * warped).astype(f32)
    n = d.size
    if kind == "l1sum":
        return float(np.abs(d.astype(np.float64)).sum()), (mask * np.sign(d)).astype(f32)
    if kind == "l2":
        return float((d.astype(np.float64) ** 2).mean()), (mask * (f32(2.0) * d) / f32(n)).astype(f32)
    if kind == "l1":
        sq = d * d + f32(1e-6)
        r = np.sqrt(sq)
        return float(r.astype(np.float64).mean()), (mask * d / r / f32(n)).astype(f32)
    raise ValueError(kind)


def psnr(img1, img2, peak=1):
    """blind_denoising.py:36-41"""
    x = (np.array(img1).squeeze() - np.array(img2).squeeze()).flatten()
    return 10 * np.log10(peak ** 2 / np.mean(x ** 2))
