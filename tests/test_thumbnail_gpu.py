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


def test_golden_juce_audiothumbnail_vectors_on_gpu(product, thumb_golden):
    """Vectors produced by the reference's own compiled juce::AudioThumbnail (level 0, both refillCache branches, the
    drawChannel mask, the jatm stream both ways), replayed on the device: rows a12 / a13 / a15 / n4 pinned."""
    from test_thumb_golden import check_against_golden
    assert check_against_golden(product, thumb_golden) >= 14


def cfg4_image():
    """BASELINE configs[3] at its real size: 1 h stereo 48 kHz (172 800 000 frames) as the 16-bit WAV file a recorder leaves
    (lib/AudioLevels.cpp:35-119): noise under a slow 0.1 Hz amplitude envelope, sparse full-scale clicks, blocks of equal
    values; seeded."""
    import thumbref as T
    n = 3600 * 48000
    rng = np.random.default_rng(4)
    data = np.empty((n, 2), np.int16)
    step = 1 << 22
    for a in range(0, n, step):
        b = min(n, a + step)
        x = rng.random((b - a, 2), np.float32) * 2.0 - 1.0
        x *= (0.55 + 0.45 * np.sin(2 * np.pi * 0.1 * np.arange(a, b, dtype=np.float64) / 48000.0)).astype(np.float32)[:, None]
        data[a:b] = np.rint(x * 32767.0).astype(np.int16)
    data[::1000003, 0] = 32767; data[500001::1000003, 1] = -32768           # clicks: the +127 / -128 clamps
    data[512 * 1000:512 * 1003] = 0                                         # equal-value fix-ups
    data[512 * 2000:512 * 2001, 0] = 32767
    hdr = T.wav_image(np.zeros((2, 0), np.float32), 48000, 16)
    import struct
    hdr[4:8] = np.frombuffer(struct.pack("<I", 36 + data.nbytes), np.uint8)
    hdr[40:44] = np.frombuffer(struct.pack("<I", data.nbytes), np.uint8)
    return np.concatenate([hdr, data.reshape(-1).view(np.uint8)]), n


def test_cfg4_full_size_one_hour_clip_bit_exact(product, oracle, juceref_optional):
    """configs[3] at the BASELINE size, bit-exact: all 337 501 x 2 level-0 pairs, 4096 pixels over the whole hour
    (82.4 thumb samples per pixel, 4096 running t += dt additions), 4096 pixels over a 10 s window (the branch that re-reads
    the source) and the 4096 x 400 column mask - product vs the CPU oracle and, where oracle/_ref travelled, vs the
    reference's own juce::AudioThumbnail run on the same file image."""
    import thumbref as T
    img, n = cfg4_image()
    dur = n / 48000.0
    windows = [(0.0, dur, 4096), (1800.0, 1810.0, 4096), (3599.5, 3600.5, 1000)]
    def run(lib):
        e = Engine(lib=lib); c = e.add_clip_wav(img); e.thumb_build(c, 512)
        r = [e.thumb_level0(c)] + [e.thumb_pixels(c, a, b, w) for a, b, w in windows] + [e.thumb_raster(c, 0, 0.0, dur, 4096, 400, 1.0)]
        e.close()
        return r
    got, exp = run(product), run(oracle)
    assert got[0].shape == (2, 337501, 2) and np.all(got[0][:, -1] == 0)
    for k, (x, y) in enumerate(zip(got, exp)):
        assert x.shape == y.shape and np.array_equal(x, y), f"product vs oracle: item {k}"
    assert got[0].min() == -127 and got[0].max() == 127      # a 16-bit file reaches -1.0 = -127 and +32767/32768 = 127; -128 needs a float file (golden case b)
    assert got[-1].sum() > 100000
    if juceref_optional is not None and hasattr(juceref_optional, "ref_thumb_create"):
        th = T.RefThumb(juceref_optional, img, 512)
        ref = [th.level0()] + [th.pixels(a, b, w) for a, b, w in windows] + [th.raster(0, 0.0, dur, 4096, 400, 1.0)]
        th.close()
        for k, (x, y) in enumerate(zip(got, ref)):
            assert np.array_equal(x, y), f"product vs juce::AudioThumbnail: item {k}"


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
