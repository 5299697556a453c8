"""File handling either side of the loop: gray read + /255 batch shape, truncating uint8 writer,
PSNR with an unclipped estimate -- blind_denoising.py:36-41,170-180,233-253.  No GPU."""
import numpy as np

from frame2frame_b200.blind_denoising import _is_tiff, _to_batch, imread_gray, imsave, psnr


def test_psnr_is_the_reference_formula_unclipped():
    rng = np.random.default_rng(0)
    ref = rng.random((1, 1, 9, 11))
    est = ref + rng.normal(0, 0.1, ref.shape)            # goes below 0 and above 1: NOT clipped before the PSNR
    x = (ref.squeeze() - est.squeeze()).flatten()
    assert psnr(ref, est) == 10 * np.log10(1.0 / np.mean(x ** 2))
    assert psnr(ref, est, peak=255) == 10 * np.log10(255 ** 2 / np.mean(x ** 2))


def test_png_round_trip_truncates_like_astype_uint8(tmp_path):
    den = np.array([[0.0, 0.5, 0.999, 1.0, 1.7, -0.3, 254.9 / 255, 127.99 / 255]])[None, None]    # [1,1,1,8]
    fn = str(tmp_path / "007.png")
    imsave(fn, np.squeeze(255. * np.clip(den, 0., 1.)).astype(np.uint8).reshape(1, 8))           # blind_denoising.py:237-238
    back = imread_gray(fn)
    assert back.dtype == np.uint8 and back.tolist() == [[0, 127, 254, 255, 255, 0, 254, 127]]   # truncation, not rounding
    b = _to_batch(back)
    assert b.shape == (1, 1, 1, 8) and b.dtype == np.float64 and b[0, 0, 0, 3] == 1.0            # [H,W] -> [1,1,H,W] / 255


def test_tiff_frames_are_float_and_unclipped(tmp_path):
    assert _is_tiff("a/%03d.tiff") and _is_tiff("x.tif") and not _is_tiff("x.png")
    den = np.array([[-3.5, 0.25, 300.75]], np.float32)
    fn = str(tmp_path / "001.tiff")
    imsave(fn, den)                                       # blind_denoising.py:234-235: 255 * estimate, no clip
    back = imread_gray(fn)
    assert back.dtype == np.float32 and back.tolist() == den.tolist()


def test_colour_input_uses_the_skimage_luma(tmp_path):
    from PIL import Image
    rgb = np.zeros((2, 2, 3), np.uint8)
    rgb[0, 0] = (255, 0, 0); rgb[0, 1] = (0, 255, 0); rgb[1, 0] = (0, 0, 255); rgb[1, 1] = (255, 255, 255)
    fn = str(tmp_path / "c.png")
    Image.fromarray(rgb).save(fn)
    g = imread_gray(fn)                                   # skimage.io.imread(as_gray=True): floats in [0,1]
    assert np.allclose(g, [[0.2125, 0.7154], [0.0721, 1.0]], atol=1e-12)
