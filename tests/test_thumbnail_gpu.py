"""WaveFormItem / juce::AudioThumbnail kernels vs the oracle: int8 min/max and the raster mask, bit-exact."""
import numpy as np
import pytest

from libzl_b200.engine import Engine
import scenarios as S

pytestmark = pytest.mark.gpu


def make_clip(n, nch=2, seed=4):
    rng = np.random.default_rng(seed)
    x = rng.uniform(-1, 1, (nch, n)).astype(np.float32)
    env = (0.55 + 0.45 * np.sin(2 * np.pi * 0.7 * np.arange(n) / 48000.0)).astype(np.float32)
    x *= env
    idx = rng.integers(0, n, max(1, n // 5000))
    x[0, idx] = 1.0; x[nch - 1, idx[::2]] = -1.0                  # full-scale clicks: +127 / -128 clamp cases
    x[:, 512 * 3:512 * 5] = 0.0                                    # equal-value fix-up
    x[0, 512 * 6:512 * 7] = 1.0
    return x


@pytest.mark.parametrize("n,nch,spt", [(512 * 700 + 301, 2, 512), (512 * 64, 1, 512), (100000, 2, 300), (511, 2, 512), (99991, 1, 77)])
def test_level0_bit_exact(product, oracle, n, nch, spt):
    pcm = make_clip(n, nch)
    out = []
    for lib in (product, oracle):
        e = Engine(lib=lib); c = e.add_clip(pcm); e.thumb_build(c, spt); out.append(e.thumb_level0(c)); e.close()
    assert out[0].shape == (nch, 1 + n // spt, 2)
    assert np.array_equal(out[0], out[1])


def test_level0_golden_read_max_levels(product, golden, oracle):
    pcm = golden["rml_pcm"]
    e = Engine(lib=product); c = e.add_clip(pcm); e.thumb_build(c, 512); got = e.thumb_level0(c); e.close()
    e = Engine(lib=oracle); c = e.add_clip(pcm); e.thumb_build(c, 512); exp = e.thumb_level0(c); e.close()
    assert np.array_equal(got, exp)


def test_pixels_and_raster_bit_exact(product, oracle):
    n = 48000 * 20 + 123
    pcm = make_clip(n)
    dur = n / 48000.0
    windows = [(0.0, dur, 4096), (0.0, dur, 777), (3.3, 17.9, 1024), (-1.0, dur + 2.0, 500),     # zoomed-out
               (5.0, 5.5, 4096), (7.0, 7.001, 300), (dur - 0.01, dur + 0.05, 256), (-0.02, 0.03, 128)]   # zoomed-in
    res = []
    for lib in (product, oracle):
        e = Engine(lib=lib); c = e.add_clip(pcm)
        r = [e.thumb_pixels(c, t0, t1, w) for (t0, t1, w) in windows]
        r += [e.thumb_raster(c, ch, 0.0, dur, 1024, 400, zoom) for ch, zoom in ((0, 1.0), (1, 2.5))]
        r += [e.thumb_raster(c, 0, 5.0, 5.5, 333, 127, 1.0)]
        assert e.thumb_pixels(c, 2.0, 1.0, 100) is None
        res.append(r); e.close()
    for k, (x, y) in enumerate(zip(*res)):
        assert x.shape == y.shape and np.array_equal(x, y), f"window {k}"
    assert res[0][-3].sum() > 0


def test_cfg4_shape_properties_large_clip(product):
    """configs[3] at a size the oracle would take too long for: size-independent properties.
    6 minutes stereo 48 kHz generated on the device side of the ABI (host upload), 4096 px."""
    n = 48000 * 360
    rng = np.random.default_rng(44)
    pcm = rng.uniform(-1, 1, (2, n)).astype(np.float32)
    pcm *= np.linspace(0.05, 1.0, n, dtype=np.float32)
    e = Engine(lib=product); c = e.add_clip(pcm)
    e.thumb_build(c, 512)
    lv = e.thumb_level0(c)
    assert lv.shape == (2, 1 + n // 512, 2) and np.all(lv[:, -1] == 0)
    blk = pcm[:, :(n // 512) * 512].reshape(2, -1, 512)
    # quantised block extrema must bracket the true extrema within one int8 step
    assert np.all(np.abs(lv[:, :-1, 1].astype(np.float32) - np.clip(np.rint(blk.max(-1) * 127.0), -128, 127)) <= 1)
    assert np.all(np.abs(lv[:, :-1, 0].astype(np.float32) - np.clip(np.rint(blk.min(-1) * 127.0), -128, 127)) <= 1)
    px = e.thumb_pixels(c, 0.0, n / 48000.0, 4096)
    # per-pixel min of mins / max of maxes: the union over all pixels equals the union over level 0
    assert px[..., 0].min() == lv[:, :-1, 0].min() and px[..., 1].max() == lv[:, :-1, 1].max()
    # idempotence
    assert np.array_equal(px, e.thumb_pixels(c, 0.0, n / 48000.0, 4096))
    e.close()
