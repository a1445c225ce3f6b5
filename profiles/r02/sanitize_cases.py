"""Small product-only runs for compute-sanitizer (memcheck / racecheck): cfg1 (Z and J), one cfg2 with every quirk and a
multi-period two-stream render with commands, a thumbnail (level 0, both pixel branches, raster), a WAV ingest + bounce."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import libzl_b200
from libzl_b200 import _capi as K
from libzl_b200.engine import Engine
import scenarios as S
import thumbref as T

p = libzl_b200.load()
a = S.cfg1_single_voice(p, mode=K.ZL_MODE_Z, looping=True, length_beats=1.5, nblocks=8)
b = S.cfg1_single_voice(p, mode=K.ZL_MODE_J, nblocks=8)
c = S.cfg2_polyphonic(p, buses=8, voices=8, nblocks=6, launches=2)
d = S.cfg2_polyphonic(p, n_sessions=2, buses=3, voices=24, nframes=256, nblocks=3)          # unstaged path
rng = np.random.default_rng(5)
pcm = rng.uniform(-1, 1, (2, 48000 * 2 + 123)).astype(np.float32)
e = Engine(lib=p)
cid = e.add_clip_wav(T.wav_image(pcm, 48000, 24))
e.thumb_build(cid, 512)
px = [e.thumb_pixels(cid, 0.0, 2.0, 512), e.thumb_pixels(cid, 0.5, 0.51, 256), e.thumb_raster(cid, 0, 0.0, 2.0, 256, 100)]
e.play(cid, midi_channel=0)
e.render(4, 512)
w = e.bounce_wav(0, 2, K.ZL_BUS_DRY, 16)
j = e.thumb_save_jatm(cid)
e.close()
print("sanitize cases ok", float(np.abs(a["raw"]).max()), float(np.abs(c["master"]).max()), w.size, j.size)
