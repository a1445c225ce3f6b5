"""Generates tests/golden/juce_thumb_golden.npz from the REFERENCE'S OWN compiled juce::AudioThumbnail
(oracle/_ref/libzl_juceref.so: juce_audio_utils/gui/juce_AudioThumbnail.cpp built in place by `make -C oracle ref`,
driven the way lib/WaveFormItem.cpp drives it; wrapper oracle/ref/juce_thumb_ref.cpp).

Run in the build container only (needs /root/reference):  python tests/golden/make_thumb_golden.py
The vectors pin rows a12 (level 0 incl. MinMaxValue::setFloat), a13 (CachedWindow::refillCache, both branches), a15
(drawChannel rectangles through the QPainter int truncation) and n4 (saveTo / loadFrom) of SURVEY.md section 8; the GPU box
replays them without the reference.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import loader  # noqa: E402
import thumbref as T  # noqa: E402


def clip(n, nch, seed):
    rng = np.random.default_rng(seed)
    x = rng.uniform(-1, 1, (nch, n)).astype(np.float32)
    x *= (0.55 + 0.45 * np.sin(2 * np.pi * 3.0 * np.arange(n) / 48000.0)).astype(np.float32)
    idx = rng.integers(0, n, max(1, n // 4000))
    x[0, idx] = 1.0; x[nch - 1, idx[::2]] = -1.0          # full-scale clicks: the +127 / -128 clamps
    x[:, 512 * 3:512 * 5] = 0.0                            # equal-value fix-up (0,0) -> (0,1)
    x[0, 512 * 6:512 * 7] = 1.0                            # (127,127) -> (126,127)
    x[nch - 1, 512 * 8:512 * 9] = -1.0
    return x


CASES = [  # name, frames, channels, rate, bits of the WAV image, seed
    ("a", 57600 + 301, 2, 48000, 16, 41),
    ("b", 24000 + 77, 1, 44100, -32, 42),
]
WINDOWS = {   # (t0, t1, width): zoomed-out and zoomed-in, offsets before the start and past the end
    "a": [(0.0, None, 512), (0.0, None, 333), (0.21, 0.93, 200), (-0.3, 1.9, 100), (0.5, 0.55, 400), (0.7, 0.7005, 64),
          (1.19, 1.25, 96), (-0.002, 0.004, 64)],
    "b": [(0.0, None, 256), (0.1, 0.12, 300)],
}
RASTERS = {"a": [(0, 0.0, None, 256, 100, 1.0), (1, 0.0, None, 200, 64, 2.5), (0, 0.5, 0.55, 128, 50, 1.0)],
           "b": [(0, 0.0, None, 128, 40, 1.0)]}


def main():
    ref = loader.load_juceref()
    if ref is None or not hasattr(ref, "ref_thumb_create"):
        loader.build(ref=True)
        ref = loader.load_juceref()
    out = {"cases": np.array([c[0] for c in CASES])}
    for name, n, nch, rate, bits, seed in CASES:
        img = T.wav_image(clip(n, nch, seed), rate, bits)
        th = T.RefThumb(ref, img, 512)
        out[f"{name}_wav"] = img
        out[f"{name}_level0"] = th.level0()
        out[f"{name}_total_length"] = np.array([th.total_length])
        wins = []
        for k, (t0, t1, w) in enumerate(WINDOWS[name]):
            t1 = th.total_length if t1 is None else t1
            wins.append((t0, t1, w))
            out[f"{name}_px_{k}"] = th.pixels(t0, t1, w)
        out[f"{name}_windows"] = np.array(wins, np.float64)
        ras = []
        for k, (ch, t0, t1, w, h, zoom) in enumerate(RASTERS[name]):
            t1 = th.total_length if t1 is None else t1
            ras.append((ch, t0, t1, w, h, zoom))
            out[f"{name}_mask_{k}"] = th.raster(ch, t0, t1, w, h, zoom)
        out[f"{name}_rasters"] = np.array(ras, np.float64)
        jatm = th.save()
        out[f"{name}_jatm"] = jatm
        back = T.ref_load_jatm(ref, jatm)
        assert back is not None and np.array_equal(back, out[f"{name}_level0"])
        th.close()
    # degenerate windows the reference refuses (refillCache returns false): nothing is drawn
    path = os.path.join(HERE, "juce_thumb_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
