"""The step after the path: HDRADC::develop() -> a multichannel OpenEXR file (hdradc.cpp:277-295) and the
scripts' dB / range-axis conversions (Receive.ipynb cells 26-27, 31)."""
import os

import numpy as np
import pytest

from beifong_b200 import configs, exr


def test_exr_round_trip_and_independent_reader(tmp_path):
    rng = np.random.default_rng(0)
    for dt in (np.float16, np.float32):
        img = rng.normal(size=(9, 6, 4)).astype(dt)
        img[0, 0, 0] = np.nan                                  # empty bins develop to NaN
        names = ["Y", "A", "phase.bin0", "phase.bin1"]
        path = exr.write_exr(str(tmp_path / f"{dt.__name__}.exr"), img, names)
        out, got = exr.read_exr(path)
        assert got == sorted(names, key=lambda s: s.encode())     # the format stores channels alphabetically
        for i, n in enumerate(names):
            assert np.array_equal(out[..., got.index(n)], img[..., i], equal_nan=True)
    # an independent OpenEXR implementation (OpenCV's bundled codec) decodes what the writer produced
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    cv2 = pytest.importorskip("cv2")
    rgb = rng.uniform(size=(5, 7, 3)).astype(np.float32)
    path = exr.write_exr(str(tmp_path / "rgb.exr"), rgb, ["R", "G", "B"])
    im = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    if im is None:
        pytest.skip("this OpenCV build has no OpenEXR codec")
    assert im.shape == (5, 7, 3) and np.array_equal(im[..., ::-1], rgb)        # OpenCV returns BGR
    with pytest.raises(ValueError):
        exr.write_exr(str(tmp_path / "bad.exr"), rgb, ["R", "G"])


@pytest.mark.gpu
def test_develop_to_exr_and_power_db(native, tmp_path):
    s = configs.small_radar(spectrum_width=1).commit()
    n = 1 << 18
    s.receive(n, seed=2)
    raw = s.profile()
    # develop(): the extension is forced to .exr, channels Y, A, phase.bin0 (weight dropped), float16 components by default
    path = s.develop(str(tmp_path / "range_profile.png"))
    assert path.endswith("range_profile.exr") and os.path.exists(path)
    img, names = exr.read_exr(path)
    assert names == ["A", "Y", "phase.bin0"] and img.dtype == np.float16 and img.shape == (raw.shape[0], raw.shape[1], 3)
    assert np.array_equal(img[..., names.index("Y")], s.bitmap(raw=False)[..., 0], equal_nan=True)
    with pytest.raises(RuntimeError):
        s.develop("")
    # Receive.ipynb cells 26-27: 10 log10(v (v > 0) / sample_count + eps) in float64
    bmp = raw[..., 0]
    want = 10 * np.log10(bmp * (bmp > 0) / n + np.finfo(float).eps)
    got = s.power_db(n, channel=0, mode="mask")
    assert got.shape == bmp.shape and np.allclose(got, want, rtol=1e-6, atol=1e-4)
    assert (got > -150).sum() > 10                             # there is signal above the eps floor
    # cell 31: 10 log10(|v| + eps)
    want = 10 * np.log10(np.abs(bmp) + np.finfo(float).eps)
    assert np.allclose(s.power_db(1, channel=0, mode="abs"), want, rtol=1e-6, atol=1e-4)
    assert np.allclose(s.range_axis(0.0, 61.44, 1024), np.linspace(0.0, 61.44, 1024) / 2)
    s.close()
