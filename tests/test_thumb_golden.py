"""Rows a12 / a13 / a15 / n4 pinned: the CPU oracle against vectors produced by the reference's own compiled
juce::AudioThumbnail (tests/golden/juce_thumb_golden.npz, generator tests/golden/make_thumb_golden.py) and, where
oracle/_ref was built, against that object live on random clips and windows.  CPU only; the same vectors are replayed on
the device in tests/test_thumbnail_gpu.py."""
import numpy as np
import pytest

from libzl_b200.engine import Engine
import thumbref as T


def check_against_golden(lib, g):
    """Runs every golden case through `lib` (oracle or product); returns the number of comparisons made."""
    done = 0
    for name in g["cases"]:
        e = Engine(lib=lib)
        c = e.add_clip_wav(g[f"{name}_wav"])
        e.thumb_build(c, 512)
        assert np.array_equal(e.thumb_level0(c), g[f"{name}_level0"]), f"{name}: level 0"
        nch = g[f"{name}_level0"].shape[0]
        rate = e.wav_probe(g[f"{name}_wav"]).sample_rate
        for k, (t0, t1, w) in enumerate(g[f"{name}_windows"]):
            got, exp = e.thumb_pixels(c, t0, t1, int(w), nch), g[f"{name}_px_{k}"]
            ok = np.ones(int(w), bool)
            if (t1 - t0) / w * rate <= 512 and t0 < 0:
                ok = T.valid_pixels(t0, t1, int(w), rate)         # the reference leaves these pixels stale; the boundary zero-fills
                assert np.all(got[:, ~ok] == 0)
            assert np.array_equal(got[:, ok], exp[:, ok]), f"{name}: window {k}"
            done += 1
        for k, (ch, t0, t1, w, h, zoom) in enumerate(g[f"{name}_rasters"]):
            got = e.thumb_raster(c, int(ch), t0, t1, int(w), int(h), float(zoom))
            assert np.array_equal(got, g[f"{name}_mask_{k}"]), f"{name}: raster {k}"
            done += 1
        assert np.array_equal(e.thumb_save_jatm(c), g[f"{name}_jatm"]), f"{name}: jatm stream"
        e2 = Engine(lib=lib)
        c2 = e2.add_clip_wav(g[f"{name}_wav"])
        e2.thumb_load_jatm(c2, g[f"{name}_jatm"])
        assert np.array_equal(e2.thumb_level0(c2), g[f"{name}_level0"]), f"{name}: jatm load"
        e.close(); e2.close()
    return done


def test_oracle_thumbnail_matches_juce_golden(oracle, thumb_golden):
    assert check_against_golden(oracle, thumb_golden) >= 14


@pytest.mark.parametrize("bits", [-32, 16, 24])
def test_oracle_thumbnail_matches_juce_live(oracle, juceref, bits):
    if not hasattr(juceref, "ref_thumb_create"):
        pytest.skip("oracle/_ref was built without the AudioThumbnail wrapper")
    rng = np.random.default_rng(100 + bits)
    for trial in range(3):
        nch = int(rng.integers(1, 3))
        n = int(rng.integers(3000, 90000))
        rate = int(rng.choice([22050, 44100, 48000, 96000]))
        x = rng.uniform(-1, 1, (nch, n)).astype(np.float32) * np.float32(rng.uniform(0.05, 1.2))
        x[:, n // 3:n // 3 + 1500] = np.float32(rng.choice([0.0, 1.0, -1.0, 0.5]))
        img = T.wav_image(x, rate, bits)
        th = T.RefThumb(juceref, img, 512)
        e = Engine(lib=oracle)
        c = e.add_clip_wav(img)
        assert np.array_equal(e.thumb_level0(c) if e.thumb_build(c, 512) is None else None, th.level0())
        dur = th.total_length
        for _ in range(6):
            w = int(rng.integers(1, 700))
            t0 = float(rng.uniform(-0.1 * dur, dur))
            t1 = t0 + float(rng.choice([rng.uniform(1e-4, 0.02), rng.uniform(0.02, 1.2 * dur)]))
            got, exp = e.thumb_pixels(c, t0, t1, w, nch), th.pixels(t0, t1, w)
            assert (got is None) == (exp is None)
            if got is None:
                continue
            ok = T.valid_pixels(t0, t1, w, rate) if ((t1 - t0) / w * rate <= 512 and t0 < 0) else np.ones(w, bool)
            assert np.array_equal(got[:, ok], exp[:, ok]), (trial, t0, t1, w)
            if t0 >= 0:
                h = int(rng.integers(1, 300)); zoom = float(rng.choice([1.0, 0.5, 3.0])); ch = int(rng.integers(0, nch))
                assert np.array_equal(e.thumb_raster(c, ch, t0, t1, w, h, zoom), th.raster(ch, t0, t1, w, h, zoom)), (trial, t0, t1, w, h)
        assert e.thumb_pixels(c, 1.0, 0.5, 10, nch) is None and th.pixels(1.0, 0.5, 10) is None
        jatm = e.thumb_save_jatm(c)
        assert np.array_equal(jatm, th.save())
        assert np.array_equal(T.ref_load_jatm(juceref, jatm), th.level0())
        th.close(); e.close()
