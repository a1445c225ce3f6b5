"""Host-logic / oracle self-consistency on CPU: quirk modes, FMA-contracted build, state machine edges."""
import numpy as np

from libzl_b200 import _capi as K
from libzl_b200.engine import Engine
import scenarios as S


def test_fma_build_within_tolerance_of_exact_build(oracle, oracle_fma):
    """The reference is built with contraction on (CMakeLists.txt:6); the exact checker with it off.
    north_star tolerance: max abs error 1e-5 of full scale."""
    a = S.cfg2_polyphonic(oracle, buses=6, nblocks=6, launches=2)
    b = S.cfg2_polyphonic(oracle_fma, buses=6, nblocks=6, launches=2)
    assert not S.compare(a, b, audio_tol=1e-5, loose_levels=True)


def test_q2_preincrement_shifts_output_by_one(oracle):
    a = S.cfg1_single_voice(oracle, quirks=K.ZL_QUIRKS_FAITHFUL, nblocks=4)
    b = S.cfg1_single_voice(oracle, quirks=K.ZL_Q1_TAP_GAIN | K.ZL_Q3_RENOTEOFF, nblocks=4)
    ra, rb = a["raw"][0, 0].reshape(4, 512), b["raw"][0, 0].reshape(4, 512)
    assert np.all(ra[:, 0] == 0.0), "index 0 of every period is never written (lib/SamplerSynthVoice.cpp:218-221)"
    assert np.array_equal(ra[:, 1:], rb[:, :-1])


def test_default_looping_clip_repeats_first_frame(oracle):
    """lengthInBeats defaults to -1 => beat-aligned branch with a zero tick length: the voice restarts
    every frame (lib/SamplerSynthVoice.cpp:227-242, ClipAudioSource.cpp:65)."""
    r = S.cfg1_single_voice(oracle, looping=True, nblocks=2)
    x = r["raw"][0, 0].reshape(2, 512)
    assert np.all(x[:, 0] == 0.0) and np.all(x[:, 2:] == x[0, 2]) and x[0, 2] != 0.0


def test_non_looping_voice_ends_and_frees_slot(oracle):
    r = S.cfg1_single_voice(oracle, nblocks=24, frames=10000)
    assert r["v_playing"][-1].sum() == 0
    # pitch 1.0: one source frame per output frame up to stopPosition = (int)(f32 seconds * rate)
    assert r["voice_frames"].sum() in (9999, 10000)


def test_voice_allocation_drops_ninth_start(oracle):
    e = Engine(lib=oracle)
    c = e.add_clip(S.noise_clip(1, 4000))
    for k in range(9):
        e.play(c, note=60 + k)
    e.render(1, 64)
    v = e.voices()
    assert sum(x.playing for x in v) == 8      # SAMPLER_CHANNEL_VOICE_COUNT, lib/SamplerSynth.cpp:23,204-215
    e.command(clip=99, midi_channel=0, start_playback=1)
    e.command(clip=c, midi_channel=50, start_playback=1)
    assert e.dropped_commands() == 2
    e.close()


def test_levels_tick_decay_and_db(oracle):
    e = Engine(lib=oracle)
    c = e.add_clip(np.full((2, 4000), 0.5, np.float32))
    e.set_clip_params(c, volume_abs=1.0)
    e.play(c, volume=1.0)
    e.render(2, 256)
    a = e.levels_tick(0, 0, 1, is_playback=True)
    assert a.peak_a > 0 and abs(a.db_a - 20 * np.log10(a.peak_a * np.float32(0.00000152587))) < 1e-4
    b = e.levels_tick(0, 0, -1, is_playback=True)
    assert b.peak_a == max(0, a.peak_a - 10000)      # lib/AudioLevels.cpp:359
    assert abs(b.hold_a - a.hold_a * np.float32(0.9)) < 1e-7 or b.hold_a == a.hold_a
    e.close()


def test_thumbnail_pixels_and_raster_edges(oracle):
    e = Engine(lib=oracle)
    n = 512 * 300 + 77
    pcm = (S.noise_clip(4, n) * np.linspace(0, 1, n, dtype=np.float32))
    c = e.add_clip(pcm)
    dur = n / 48000.0
    px = e.thumb_pixels(c, 0.0, dur, 128)
    assert px.shape == (2, 128, 2)
    assert np.all(px[..., 1] >= px[..., 0])
    assert e.thumb_pixels(c, 1.0, 1.0, 128) is None          # timePerPixel <= 0 => refillCache false
    assert e.thumb_pixels(c, 0.0, 1.0, 0) is None
    zoom = e.thumb_pixels(c, 1.0, 1.01, 256)                  # zoomed-in branch (direct read)
    assert zoom.shape == (2, 256, 2)
    past = e.thumb_pixels(c, dur + 1.0, dur + 1.001, 16)      # past the end => (0,0)
    assert np.all(past == 0)
    m = e.thumb_raster(c, 0, 0.0, dur, 128, 64)
    assert m.shape == (64, 128) and m.max() == 1
    cols = m.sum(axis=0)
    assert cols[-1] >= cols[1]                                # louder at the end
    e.close()
