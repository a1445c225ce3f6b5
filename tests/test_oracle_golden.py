"""The CPU oracle (oracle/zl_oracle.cpp) pinned against golden vectors generated from the reference's
own compiled JUCE objects (tests/golden/make_golden.py) and, where oracle/_ref exists, against those
objects live.  CPU only."""
import ctypes as C

import numpy as np
import pytest

from libzl_b200 import _capi as K
from libzl_b200.engine import Engine

f32p = C.POINTER(C.c_float)


def fp(a):
    return a.ctypes.data_as(f32p)


def test_adsr_matches_juce_golden(oracle, golden):
    params = golden["adsr_params"]
    for i, (sr, a, d, s, r, n_on, n_total, ren) in enumerate(params):
        buf = np.zeros(int(n_total), np.float32)
        oracle.dll.zlo_adsr_run(sr, a, d, s, r, int(n_on), int(n_total), int(ren), fp(buf))
        assert np.array_equal(buf, golden[f"adsr_{i}"]), f"ADSR case {i} differs from juce::ADSR"


def test_adsr_matches_juce_live(oracle, juceref):
    rng = np.random.default_rng(5)
    for _ in range(40):
        sr = float(rng.choice([22050.0, 44100.0, 48000.0, 96000.0]))
        a, d, r = (float(np.float32(rng.choice([0.0, rng.uniform(0, 0.05)]))) for _ in range(3))
        s = float(np.float32(rng.uniform(0.1, 1.0)))
        n_total = 3000
        n_on = int(rng.integers(1, 2500))
        ren = int(rng.integers(0, 2))
        x, y = np.zeros(n_total, np.float32), np.zeros(n_total, np.float32)
        oracle.dll.zlo_adsr_run(sr, a, d, s, r, n_on, n_total, ren, fp(x))
        juceref.ref_adsr_run(sr, a, d, s, r, n_on, n_total, ren, fp(y))
        assert np.array_equal(x, y)


def _adsr(oracle, sr, a, d, s, r, n_on, n_total):
    buf = np.zeros(n_total, np.float32)
    oracle.dll.zlo_adsr_run(sr, a, d, s, r, n_on, n_total, 0, fp(buf))
    return buf


def test_adsr_properties_of_juce_unit_tests(oracle):
    """The assertions of ADSRTests (juce_ADSR_test.cpp:26-243) on the restated envelope."""
    sr, a, d, s, r = 44100.0, 0.1, 0.1, 0.5, 0.1
    na, nd = int(a * sr), int(d * sr)
    env = _adsr(oracle, sr, a, d, s, r, na + nd + 5000, na + nd + 5000 + int(r * sr) + 100)
    att = env[:na - 1]
    assert np.all(np.diff(att) > 0), "attack strictly increasing"
    dec = env[na + 1:na + nd - 1]
    assert np.all(np.diff(dec) < 0), "decay strictly decreasing"
    sus = env[na + nd + 10:na + nd + 4000]
    assert sus.min() == sus.max() == np.float32(s), "sustain exact"
    rel = env[na + nd + 5000:na + nd + 5000 + int(r * sr) - 2]
    assert np.all(np.diff(rel) < 0), "release strictly decreasing"
    assert env[-1] == 0.0, "idle after release"
    # zero-length attack jumps to decay at 1.0; zero-length decay jumps to sustain; zero release -> idle
    e = _adsr(oracle, sr, 0.0, d, s, r, 100, 200)
    assert e[0] < 1.0 and e[0] > 0.99
    e = _adsr(oracle, sr, a, 0.0, s, r, na + 50, na + 60)
    assert e[na + 5] == np.float32(s)
    e = _adsr(oracle, sr, a, d, s, 0.0, 1000, 1100)
    assert e[1000] == 0.0 and e[1050] == 0.0


def check_adsr_golden_through_voice(lib, golden):
    """Every golden juce::ADSR sequence, replayed through the voice path of `lib` (J-mode; the renote cases through
    Z-mode quirk Q3)."""
    import scenarios as S
    for i, (sr, a, d, s, r, n_on, n_total, ren) in enumerate(golden["adsr_params"]):
        env, n = S.adsr_via_voice(lib, sr, a, d, s, r, n_on, n_total, renote=bool(ren))
        assert n > n_on
        assert np.array_equal(env[:n], golden[f"adsr_{i}"][:n]), f"ADSR case {i} through the voice differs from juce::ADSR"


def test_adsr_golden_through_the_oracle_voice(oracle, golden):
    check_adsr_golden_through_voice(oracle, golden)


def check_adsr_properties_through_voice(lib):
    """The assertions of ADSRTests (juce_ADSR_test.cpp:26-243) on the envelope a voice of `lib` produces."""
    import scenarios as S
    sr, a, d, s, r = 44100.0, 0.1, 0.1, 0.5, 0.1
    na, nd = int(a * sr), int(d * sr)
    n_on = ((na + nd + 5000) // 100) * 100
    n_total = ((n_on + int(r * sr) + 300) // 100) * 100
    env, _ = S.adsr_via_voice(lib, sr, a, d, s, r, n_on, n_total)
    assert np.all(np.diff(env[:na - 1]) > 0), "attack strictly increasing"
    assert np.all(np.diff(env[na + 1:na + nd - 1]) < 0), "decay strictly decreasing"
    sus = env[na + nd + 10:n_on - 10]
    assert sus.min() == sus.max() == np.float32(s), "sustain exact"
    assert np.all(np.diff(env[n_on:n_on + int(r * sr) - 2]) < 0), "release strictly decreasing"
    assert env[-1] == 0.0, "idle after release"
    e, _ = S.adsr_via_voice(lib, sr, 0.0, d, s, r, 100, 200)
    assert e[0] < 1.0 and e[0] > 0.99                                   # zero-length attack: decay starts from 1.0
    e, _ = S.adsr_via_voice(lib, sr, a, 0.0, s, r, ((na + 100) // 100) * 100, ((na + 300) // 100) * 100)
    assert e[na + 5] == np.float32(s)                                   # zero-length decay: straight to sustain
    e, _ = S.adsr_via_voice(lib, sr, a, d, s, 0.0, 1000, 1100)
    assert e[1000] == 0.0 and e[1050] == 0.0                            # zero-length release: idle at once


def test_adsr_properties_through_the_oracle_voice(oracle):
    check_adsr_properties_through_voice(oracle)


def _run_jmode(lib, p, pcm):
    nch, n, rate, orate, root, note, vel, att, rel, nb, nf, off = p
    e = Engine(lib=lib, mode=K.ZL_MODE_J, sample_rate=float(orate), max_frames=1024, max_blocks=64)
    c = e.add_clip(pcm, rate=float(rate))
    e.set_clip_params(c, attack=float(np.float32(att)), decay=0.1, sustain=1.0, release=float(np.float32(rel)), root_note=int(root))
    e.play(c, midi_channel=-2, note=int(note), volume=float(np.float32(vel)))
    if off >= 0:
        e.stop(c, midi_channel=-2, note=int(note), block=int(off))
    e.render(int(nb), int(nf))
    l, r = e.bus(0, 0, K.ZL_BUS_RAW)
    e.close()
    return l, r


def test_jmode_voice_matches_juce_golden(oracle, golden):
    for i, p in enumerate(golden["j_params"]):
        l, r = _run_jmode(oracle, p, golden[f"j_pcm_{i}"])
        assert np.array_equal(l, golden[f"j_L_{i}"]), f"J-mode case {i} L differs from juce::SamplerVoice::renderNextBlock"
        assert np.array_equal(r, golden[f"j_R_{i}"]), f"J-mode case {i} R differs"


def test_jmode_voice_matches_juce_live(oracle, juceref):
    rng = np.random.default_rng(11)
    for _ in range(6):
        nch = int(rng.integers(1, 3)); n = int(rng.integers(500, 5000))
        p = (nch, n, float(rng.choice([44100.0, 48000.0])), 48000.0, 60, int(rng.integers(48, 73)), float(np.float32(rng.uniform(0.2, 1))),
             float(np.float32(rng.uniform(0, 0.02))), float(np.float32(rng.uniform(0.005, 0.05))), 6, 256, int(rng.integers(-1, 5)))
        pcm = rng.uniform(-1, 1, (nch, n)).astype(np.float32)
        L = np.zeros(6 * 256, np.float32); R = np.zeros(6 * 256, np.float32)
        ptrs = (f32p * nch)(*[fp(pcm[k]) for k in range(nch)])
        juceref.ref_jvoice_render(ptrs, nch, n, p[2], p[3], p[4], p[5], p[6], p[7], p[8], p[9], p[10], p[11], fp(L), fp(R))
        l, r = _run_jmode(oracle, p, pcm)
        assert np.array_equal(l, L) and np.array_equal(r, R)


def test_thumbnail_level0_matches_read_max_levels_golden(oracle, golden):
    pcm, lv = golden["rml_pcm"], golden["rml_levels"]
    e = Engine(lib=oracle)
    c = e.add_clip(pcm)
    e.thumb_build(c, 512)
    got = e.thumb_level0(c)
    assert got.shape == (2, 1 + pcm.shape[1] // 512, 2)
    exp = np.zeros((2, 40, 2), np.int8)
    for b in range(40):
        for ch in range(2):
            o = (C.c_int8 * 2)()
            oracle.dll.zlo_set_float(float(lv[b, 2 * ch]), float(lv[b, 2 * ch + 1]), o)
            exp[ch, b] = (o[0], o[1])
    assert np.array_equal(got[:, :40], exp)
    assert np.array_equal(got[:, 40], np.zeros((2, 2), np.int8)), "trailing thumb sample stays (0,0)"
    # equal-value fix-up and clamp cases are present
    assert (exp[0, 7] == (32, 33)).all() and (exp[1, 9] == (0, 1)).all()
    e.close()


def test_round_to_int_matches_juce_golden(oracle, golden):
    got = np.array([oracle.dll.zlo_round_to_int(float(x)) for x in golden["rti_x"]], np.int32)
    assert np.array_equal(got, golden["rti_d"])
    gotf = np.array([oracle.dll.zlo_round_to_int(float(np.float32(x))) for x in golden["rti_x"]], np.int32)
    assert np.array_equal(gotf, golden["rti_f"])


def test_rms_and_magnitude_match_juce_golden(oracle, golden):
    x = golden["rms_x"]
    e = Engine(lib=oracle)
    pa, pb, sa, sb = e.levels_scan(x, -x)
    assert np.float32(np.sqrt(sa / x.size)) == golden["rms"][0]
    assert sa == sb and pa == pb
    assert pa == int(np.float32(131072.0) * golden["rms"][1])      # peak of |x| is getMagnitude
    e.close()
