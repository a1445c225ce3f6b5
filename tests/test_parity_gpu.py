"""Parity of the CUDA product (through the C ABI) against the CPU oracle on the same seeded inputs.

Bar (BASELINE.json north_star): float audio within max abs 1e-5 of full scale; thumbnail min/max, integer
peak meters, voice state (positions, playing flags, frame counts) bit-exact.  The kernels use explicit
round-to-nearest arithmetic in the reference's evaluation order, so against the -ffp-contract=off oracle
the audio is in fact required to be bit-identical here; the FMA-contracted oracle build is held to 1e-5.
"""
import numpy as np
import pytest

from libzl_b200 import _capi as K
from libzl_b200.engine import Engine
import scenarios as S

pytestmark = pytest.mark.gpu

AUDIO_TOL = 1e-5   # north_star: max abs error 1e-5 of full scale


def both(fn, product, oracle, **kw):
    return fn(product, **kw), fn(oracle, **kw)


@pytest.mark.parametrize("mode", [K.ZL_MODE_Z, K.ZL_MODE_J])
@pytest.mark.parametrize("looping", [False, True])
def test_cfg1_single_voice(product, oracle, mode, looping):
    a, b = both(S.cfg1_single_voice, product, oracle, mode=mode, looping=looping, length_beats=(1.5 if looping else -1.0))
    assert not S.compare(a, b, audio_tol=AUDIO_TOL, exact_audio=True)
    assert a["voice_frames"].sum() > 0


def test_cfg1_default_looping_restart_every_frame(product, oracle):
    a, b = both(S.cfg1_single_voice, product, oracle, looping=True, nblocks=3)
    assert not S.compare(a, b, exact_audio=True)


@pytest.mark.parametrize("quirks", [K.ZL_QUIRKS_FAITHFUL, 0, K.ZL_Q1_TAP_GAIN, K.ZL_Q2_PREINC, K.ZL_Q3_RENOTEOFF])
def test_cfg2_polyphonic_quirk_modes(product, oracle, quirks):
    a, b = both(S.cfg2_polyphonic, product, oracle, quirks=quirks, buses=16, voices=8)
    assert not S.compare(a, b, audio_tol=AUDIO_TOL, exact_audio=True)
    assert a["voice_frames"].sum() > 100000


def test_cfg2_jmode(product, oracle):
    a, b = both(S.cfg2_polyphonic, product, oracle, mode=K.ZL_MODE_J, buses=16, voices=8)
    assert not S.compare(a, b, audio_tol=AUDIO_TOL, exact_audio=True)


def test_cfg2_against_fma_contracted_reference_build(product, oracle_fma):
    a, b = both(S.cfg2_polyphonic, product, oracle_fma, buses=8, voices=8)
    assert not S.compare(a, b, audio_tol=AUDIO_TOL, loose_levels=True)


def test_cfg2_multi_session_and_odd_shapes(product, oracle):
    a, b = both(S.cfg2_polyphonic, product, oracle, n_sessions=5, buses=3, voices=4, nframes=300, nblocks=7)
    assert not S.compare(a, b, exact_audio=True)
    a, b = both(S.cfg2_polyphonic, product, oracle, n_sessions=2, buses=12, voices=8, nframes=1024, nblocks=3, clip_frames=9000)
    assert not S.compare(a, b, exact_audio=True)
    a, b = both(S.cfg2_polyphonic, product, oracle, n_sessions=2, buses=5, voices=16, nframes=256, nblocks=5)
    assert not S.compare(a, b, exact_audio=True)
    a, b = both(S.cfg2_polyphonic, product, oracle, n_sessions=3, buses=4, voices=2, nframes=1, nblocks=9)
    assert not S.compare(a, b, exact_audio=True)


def test_cfg2_master_only_outputs(product, oracle):
    outs = K.ZL_OUT_MASTER | K.ZL_OUT_LEVELS
    a, b = both(S.cfg2_polyphonic, product, oracle, n_sessions=4, buses=8, outputs=outs)
    assert "raw" not in a and "master" in a
    assert not S.compare(a, b, exact_audio=True)


@pytest.mark.parametrize("beat_aligned", [True, False])
def test_cfg3_sketch_looping_tracks_with_meters(product, oracle, beat_aligned):
    a, b = both(S.cfg3_sketch, product, oracle, beat_aligned=beat_aligned)
    assert not S.compare(a, b, audio_tol=AUDIO_TOL, exact_audio=True)
    assert a["meter_peaks"].max() > 0


def test_disabled_bus_and_muted_passthrough(product, oracle):
    def run(lib):
        e = Engine(lib=lib)
        c = e.add_clip(S.noise_clip(9, 5000))
        e.play(c, midi_channel=0); e.play(c, midi_channel=1, note=64); e.play(c, midi_channel=2, note=57)
        e.set_bus_enabled(0, 2, False)
        e.set_passthrough(0, 3, muted=True)
        e.set_passthrough(0, 4, dry=0.0, fx1=1.0, fx2=0.5, pan=0.0)
        e.render(3, 256)
        r = S.collect(e)
        e.close()
        return r
    a, b = run(product), run(oracle)
    assert not S.compare(a, b, exact_audio=True)
    assert np.all(a["raw"][2] == 0) and np.all(a["dry"][3] == 0) and np.any(a["raw"][3] != 0)


def test_standalone_passthrough_and_levels(product, oracle):
    rng = np.random.default_rng(12)
    for n in (1, 255, 4096, 100003):
        l = rng.uniform(-1.5, 1.5, n).astype(np.float32); r = rng.uniform(-1.5, 1.5, n).astype(np.float32)
        eg, eo = Engine(lib=product), Engine(lib=oracle)
        for kw in (dict(), dict(dry=0.8, fx1=0.3, fx2=0.0, pan=0.25), dict(pan=-0.7), dict(muted=True), dict(dry=0.0, fx1=1.0, fx2=0.5)):
            assert np.array_equal(eg.passthrough_process(l, r, **kw), eo.passthrough_process(l, r, **kw))
        pa, pb, sa, sb = eg.levels_scan(l, r)
        qa, qb, ta, tb = eo.levels_scan(l, r)
        assert (pa, pb) == (qa, qb)                                   # integer peaks: bit-exact
        assert abs(sa - ta) <= 1e-9 * ta and abs(sb - tb) <= 1e-9 * tb   # f64 sum, order-free
        eg.close(); eo.close()


def test_levels_tick_sequence_bit_exact(product, oracle):
    def run(lib):
        e = Engine(lib=lib)
        c = e.add_clip(S.sweep_clip(20000))
        e.set_clip_params(c, volume_abs=1.0, release=0.1)
        e.play(c, midi_channel=-2, volume=0.9)
        seq = []
        for k in range(12):
            e.render(4, 512, e.make_clocks(4, 512, first_block=4 * k))
            x = e.levels_tick(0, 0, 3 if k % 3 else -1, is_playback=True)
            seq.append((x.peak_a, x.peak_b, x.db_a, x.db_b, x.db_sum, x.hold_a, x.hold_b, x.rms_a, x.rms_b))
        e.close()
        return np.array(seq)
    a, b = run(product), run(oracle)
    assert np.array_equal(a[:, :2], b[:, :2])
    assert np.allclose(a[:, 2:], b[:, 2:], rtol=1e-6, atol=1e-4)      # dB through libm log10f: 1e-4 dB


def test_golden_jmode_vectors_on_gpu(product, golden):
    """The committed vectors from the real juce::SamplerVoice::renderNextBlock, replayed on the device."""
    from test_oracle_golden import _run_jmode
    for i, p in enumerate(golden["j_params"]):
        l, r = _run_jmode(product, p, golden[f"j_pcm_{i}"])
        assert np.array_equal(l, golden[f"j_L_{i}"]) and np.array_equal(r, golden[f"j_R_{i}"]), i


def test_golden_adsr_sequences_on_gpu(product, golden):
    """The committed juce::ADSR vectors (from the reference's own compiled object) replayed through the DEVICE voice:
    J-mode for the plain sequences, Z-mode quirk Q3 for the re-armed noteOff ones."""
    from test_oracle_golden import check_adsr_golden_through_voice
    check_adsr_golden_through_voice(product, golden)


def test_adsr_unit_test_properties_on_the_device_envelope(product):
    """ADSRTests (juce_ADSR_test.cpp:26-243) ported: monotone attack / decay / release, exact sustain, zero-length
    segments - on the envelope the device voice produces."""
    from test_oracle_golden import check_adsr_properties_through_voice
    check_adsr_properties_through_voice(product)


def test_unsynced_one_period_render_then_multi_period_render(product, oracle):
    """A one-stream render (one period) immediately followed - no host sync - by a two-stream render: the second render's
    counter resets and command upload (planner stream) must wait for the first (context stream)."""
    def run(lib):
        rng = np.random.default_rng(33)
        e = Engine(lib=lib, n_sessions=64, buses_per_session=4, voices_per_bus=8, max_frames=512, max_blocks=8)
        clips = [e.add_clip(rng.uniform(-1, 1, (2, 30000)).astype(np.float32)) for _ in range(3)]
        for c in clips:
            e.set_clip_params(c, attack=0.004, decay=0.02, sustain=0.6, release=0.03, pan=-0.3, length_beats=0.75)
        for s in range(64):
            for k in range(16):
                e.play(clips[k % 3], loop=bool(k % 3 == 0), midi_channel=k % 4 - 2, note=50 + k + s % 7, volume=0.7, session=s)
        outs = []
        first = 0
        for rep in range(6):
            for s in range(0, 64, 5):
                e.play(clips[rep % 3], midi_channel=rep % 4 - 2, note=70 + rep, volume=0.5, session=s, block=0)
            e.render(1, 512, e.make_clocks(1, 512, first_block=first)); first += 1          # one stream
            m1 = None
            for s in range(0, 64, 7):
                e.stop(clips[rep % 3], midi_channel=rep % 4 - 2, note=70 + rep, session=s, block=1)
            e.render(6, 512, e.make_clocks(6, 512, first_block=first)); first += 6          # two streams, queued right behind
            outs.append((e.master_all().copy(), e.block_levels()[0].copy(), e.voice_frames()))
        e.close()
        return outs
    for (m0, p0, v0), (m1, p1, v1) in zip(run(product), run(oracle)):
        assert np.array_equal(m0, m1) and np.array_equal(p0, p1) and v0 == v1
        assert np.abs(m0).max() > 0.05


def test_destroying_a_clip_in_use_stops_its_voices(product, oracle):
    """zl_clip_destroy while voices still play the clip: they end as by stopNote(0, false) (both modes), the slots are
    free again, other clips keep sounding, and nothing reads the freed PCM."""
    for mode in (K.ZL_MODE_Z, K.ZL_MODE_J):
        def run(lib):
            rng = np.random.default_rng(8)
            e = Engine(lib=lib, mode=mode, n_sessions=2, buses_per_session=2, voices_per_bus=4)
            a = e.add_clip(rng.uniform(-1, 1, (2, 20000)).astype(np.float32))
            b = e.add_clip(rng.uniform(-1, 1, (2, 20000)).astype(np.float32))
            for s in range(2):
                for k in range(4):
                    e.play(a if k % 2 else b, loop=bool(k == 0), midi_channel=k % 2 - 2, note=60 + k, volume=0.6, session=s)
            e.render(2, 256)
            r0 = S.collect(e)
            lib.clip_destroy(e.ctx, a)
            e.render(2, 256, e.make_clocks(2, 256, first_block=2))
            r1 = S.collect(e)
            for s in range(2):                                   # the freed slots take new voices
                e.play(b, midi_channel=-1, note=72, volume=0.4, session=s)
                e.play(b, midi_channel=-1, note=74, volume=0.4, session=s)
            e.render(2, 256, e.make_clocks(2, 256, first_block=4))
            r2 = S.collect(e)
            e.close()
            return r0, r1, r2
        for x, y in zip(run(product), run(oracle)):
            assert not S.compare(x, y, exact_audio=True)
        assert np.all(run(product)[1]["v_clip"] != 0) and np.any(run(product)[1]["v_clip"] == 1)


def _reference_scenarios():
    import libref as R
    return sorted(R.SCENARIOS)


@pytest.mark.parametrize("sched_batch", [1, 8])
@pytest.mark.parametrize("name", _reference_scenarios())
def test_device_matches_the_reference_compiled_hot_path(product, name, sched_batch):
    """The product against the reference's OWN compiled SamplerSynthVoice / SamplerChannel / JackPassthrough / positions model
    / AudioLevels / SyncTimer (oracle/ref/libzl_ref.cpp, synctimer_ref.cpp): through the committed digests of the reference's
    outputs and, where oracle/_ref travelled to this box, against the reference run live - bit for bit (audio planes, integer
    meter peaks, positions; for the scheduler scenarios also every period's clock and emitted command).  sched_batch 8: the
    scheduler advances eight periods per call and the device renders them as one batch (row n2's point)."""
    import json, os
    import libref as R
    golden = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "libzl_digests.json")))
    ops = R.SCENARIOS[name]()
    if sched_batch != 1 and not name.startswith("scheduler"):
        pytest.skip("only the scheduler scenarios batch periods")
    got = R.run_engine(product, ops, sched_batch=sched_batch)
    if R.available():
        assert not R.compare(R.run_reference(ops), got, skip_buses=R.disabled_buses(ops))
    assert R.digest(got, R.disabled_buses(ops)) == golden[name]


def test_native_library_is_what_ran(product):
    e = Engine(lib=product)
    c = e.add_clip(S.noise_clip(1, 2000))
    e.play(c)
    before = e.launch_count()
    e.render(1, 128)
    assert e.launch_count() >= before + 1 and e.voice_frames() == 128   # render_kernel (+ master_mix_kernel)
    import os
    import libzl_b200
    maps = open("/proc/self/maps").read()
    assert os.path.basename(libzl_b200.LIB_PATH) in maps          # ZL_B200_LIB: a kernel-variant build runs the whole suite too
    e.close()


def test_pipelined_master_readback_matches_synchronous(product, oracle):
    """zl_read_master_all_async + zl_read_wait (two master buffers, copy overlapping the next render) must deliver
    exactly what the synchronous read-back and the oracle deliver, render after render."""
    import ctypes as C

    def run(lib, pipelined):
        rng = np.random.default_rng(11)
        e = Engine(lib=lib, n_sessions=3, buses_per_session=2, voices_per_bus=4)
        c = e.add_clip(rng.uniform(-1, 1, (2, 30000)).astype(np.float32))
        e.set_clip_params(c, attack=0.002, decay=0.02, sustain=0.6, release=0.03, pan=-0.2, length_beats=1.5)
        for s in range(3):
            for k in range(4):
                e.play(c, loop=bool(k & 1), midi_channel=k % 2 - 2, note=57 + 3 * k + s, volume=0.7, session=s)
        n = 2 * 256
        outs = []
        bufs = []
        if pipelined:
            for _ in range(2):
                p = C.c_void_p()
                e._chk(lib.host_alloc(C.byref(p), 3 * 2 * n * 4), "host_alloc")
                bufs.append(p)
        for step in range(5):
            e.render(2, 256, e.make_clocks(2, 256, 48000.0, first_block=2 * step))
            if pipelined:
                hp = C.cast(bufs[step & 1], C.POINTER(C.c_float))
                e._chk(lib.read_master_all_async(e.ctx, hp, n), "read_master_all_async")
                if step == 4:
                    e._chk(lib.read_wait(e.ctx), "read_wait")
            else:
                outs.append(e.master_all().copy())
        if pipelined:
            # buffers hold steps 3 and 4 after the final wait
            res = [np.ctypeslib.as_array(C.cast(bufs[k], C.POINTER(C.c_float)), shape=(3, 2, n)).copy() for k in (1, 0)]
            for p in bufs:
                lib.host_free(p)
            e.close()
            return res            # [step 3, step 4]
        e.close()
        return outs[3:5]

    sync = run(product, False)
    pipe = run(product, True)
    ref = run(oracle, False)
    assert np.abs(sync[0]).max() > 0.05
    for a, b, c in zip(sync, pipe, ref):
        assert np.array_equal(a, b)
        assert np.array_equal(a, c)


def test_wide_buses_take_the_unstaged_and_atomic_paths(product, oracle):
    """V = 24 (> 16: descriptors straight from L2, no window staging) and V = 40 (> 32: positions-model gain by atomicMax,
    single-stream render) must match the oracle exactly like the common V <= 16 layout."""
    for voices in (24, 40):
        a, b = both(S.cfg2_polyphonic, product, oracle, n_sessions=2, buses=2, voices=voices, nframes=256, nblocks=6)
        assert not S.compare(a, b, exact_audio=True)
        assert a["voice_frames"].sum() > 10000


def test_high_pitch_ratios_bypass_the_window_staging(product, oracle):
    """Ratios above 2.4 (source windows larger than a staging slot: 96 kHz clips played two octaves up, ratio up to 8)
    and below 0.1 take the direct-gather path; mixed on one bus with staged voices the slot order must hold."""
    def run(lib):
        rng = np.random.default_rng(21)
        e = Engine(lib=lib, n_sessions=2, buses_per_session=2, voices_per_bus=8)
        clips = []
        for k in range(4):
            pcm = rng.uniform(-1, 1, (2, 60000)).astype(np.float32)
            c = e.add_clip(pcm, 96000.0 if k & 1 else 48000.0)
            e.set_clip_params(c, attack=0.001, decay=0.02, sustain=0.7, release=0.05, pan=float(rng.uniform(-1, 1)),
                              length_beats=float(rng.choice([1.5, 2.25])), root_note=60)
            clips.append(c)
        for s in range(2):
            for b in range(2):
                for v in range(8):
                    e.play(clips[(v + b) % 4], loop=bool(v & 1), midi_channel=b - 2, note=int(rng.choice([24, 36, 60, 72, 84])),
                           volume=float(rng.uniform(0.3, 1.0)), session=s)
        out = []
        for step in range(3):
            e.render(4, 512, e.make_clocks(4, 512, first_block=4 * step))
            out.append((e.master_all().copy(), np.stack([np.stack(e.bus(s, b, K.ZL_BUS_RAW)) for s in range(2) for b in range(2)]),
                        e.block_levels()[0].copy(), e.voice_frames()))
        e.close()
        return out
    for (m0, r0, p0, v0), (m1, r1, p1, v1) in zip(run(product), run(oracle)):
        assert np.array_equal(m0, m1) and np.array_equal(r0, r1) and np.array_equal(p0, p1) and v0 == v1
        assert np.abs(m0).max() > 0.05


def test_stream_schedules_agree_render_after_render(product, oracle):
    """The two-stream schedule (planner of the next period beside the gather, planner of period 0 chained to the previous
    render, master mix sliced by period on a third stream, uploads on the planner stream) must give bit for bit what the
    one-stream schedule and the oracle give - across renders with commands in between, a change of period size (the
    hand-over buffers move: the chain must fall back) and a one-period render in the middle."""
    def run(lib, overlap):
        rng = np.random.default_rng(21)
        e = Engine(lib=lib, n_sessions=2, buses_per_session=3, voices_per_bus=4, max_frames=512, max_blocks=8)
        if lib is product:
            e._chk(lib.ctx_set_overlap(e.ctx, 1 if overlap else 0), "ctx_set_overlap")
        clips = [e.add_clip(rng.uniform(-1, 1, (2, 9000 + 1000 * k)).astype(np.float32)) for k in range(3)]
        for c in clips:
            e.set_clip_params(c, attack=0.003, decay=0.01, sustain=0.7, release=0.02, pan=0.3, length_beats=0.75)
        for s in range(2):
            for k in range(6):
                e.play(clips[k % 3], loop=bool(k % 2), midi_channel=k % 3 - 2, note=55 + 2 * k + s, volume=0.8, session=s)
        acc = {}
        first = 0
        for step, (nb, nf) in enumerate([(4, 256), (4, 256), (3, 256), (1, 256), (5, 128), (5, 128), (2, 512), (4, 256)]):
            if step in (2, 5):
                for s in range(2):
                    e.stop(clips[step % 3], midi_channel=-2, note=55 + s, session=s, block=1 % nb)
                    e.play(clips[(step + 1) % 3], midi_channel=-1, note=64, volume=0.5, session=s, block=nb - 1)
            e.render(nb, nf, e.make_clocks(nb, nf, 48000.0, first_block=first))
            first += nb
            S.append(acc, S.collect(e))
        e.close()
        return S.finish(acc)

    two, one, ref = run(product, True), run(product, False), run(oracle, False)
    assert np.abs(two["master"]).max() > 0.05
    assert not S.compare(two, one, exact_audio=True)
    assert not S.compare(two, ref, exact_audio=True)


def test_command_ring_overflow_and_late_commands_are_dropped_like_the_reference(product, oracle):
    """lib/SamplerSynth.cpp:32,332-334: the per-bus command ring holds 256 commands per period; what does not fit is
    dropped, and so are commands addressed to a bus that does not exist or to a period this render does not have.
    A burst of 300 commands on one bus in one period takes the library's sorted slow path; the same burst split over
    two periods (150 + 150) must lose nothing on the fast path."""
    def run(lib, split):
        rng = np.random.default_rng(5)
        e = Engine(lib=lib, n_sessions=2, buses_per_session=3, voices_per_bus=8)
        c = e.add_clip(rng.uniform(-1, 1, (2, 20000)).astype(np.float32))
        e.set_clip_params(c, attack=0.0, decay=0.0, sustain=1.0, release=0.01)
        for k in range(300):                         # start / stop pairs and volume updates on session 1, bus 0
            blk = (k // 150) if split else 0
            if k % 3 == 0:
                e.play(c, midi_channel=-2, note=48 + (k // 3) % 24, volume=0.3 + 0.002 * k, session=1, block=blk)
            elif k % 3 == 1:
                e.stop(c, midi_channel=-2, note=48 + (k // 3) % 24, session=1, block=blk)
            else:
                e.play(c, midi_channel=-2, note=72, volume=0.5, session=1, block=blk)
        e.play(c, midi_channel=-1, note=60, session=0, block=2)          # survives: another bus, another period
        e.play(c, midi_channel=-2, note=61, session=0, block=3)          # period 3 of a 3-period render: dropped
        e.play(c, midi_channel=5, note=62, session=0, block=0)           # bus 7 of 3: dropped on arrival
        e.render(3, 256, e.make_clocks(3, 256))
        out = S.collect(e)
        out["dropped"] = np.array([e.dropped_commands()])
        e.close()
        return out

    for split in (False, True):
        a, b = run(product, split), run(oracle, split)
        assert not S.compare(a, b, exact_audio=True)
        assert int(a["dropped"][0]) == (300 - 256 + 2 if not split else 2)
        assert np.abs(a["raw"]).max() > 0.01


def test_live_period_as_a_graph_launch_with_a_knob_turned_every_period(product, oracle):
    """nblocks == 1 renders go out as one CUDA-graph launch per command slot (re-captured when an argument changes), and a
    passthrough / clip parameter changed between periods reaches the device through the double-buffered pinned tables
    without a stream synchronise: 40 periods with a pan knob moved every period, commands in some, a clip parameter and the
    frame count changed mid-way - graph on, graph off and the oracle give the same bytes."""
    outs = []
    for lib, graph in ((product, 1), (product, 0), (oracle, 1)):
        e = Engine(lib=lib, mode=K.ZL_MODE_Z, quirks=K.ZL_QUIRKS_FAITHFUL, max_frames=256, max_blocks=4)
        e._chk(lib.ctx_set_graph(e.ctx, graph), "ctx_set_graph")
        a = e.add_clip(S.noise_clip(3, 9000))
        b = e.add_clip(S.noise_clip(4, 30000))
        e.set_clip_params(b, length_s=0.25, length_beats=0.5, attack=0.01, release=0.05)
        e.play(a, note=60, volume=0.9)
        res = []
        for i in range(40):
            nf = 256 if i < 25 else 192
            e.set_passthrough(0, 3, dry=0.5 + 0.01 * i, fx1=0.3, fx2=0.1, pan=-0.8 + 0.04 * i)
            if i % 7 == 3:
                e.play(b, loop=True, note=57 + i % 5, volume=0.7, midi_channel=1)
            if i == 20:
                e.set_clip_params(b, pan=0.5)
            if i == 30:
                e.render(2, nf, e.make_clocks(2, nf, first_block=1000))      # a two-period render in between: the plain path
                res.append(S.collect(e))
            e.render(1, nf, e.make_clocks(1, nf, first_block=i))
            res.append(S.collect(e))
        outs.append(res)
        e.close()
    for k in (1, 2):
        for x, y in zip(outs[0], outs[k]):
            assert not S.compare(x, y, exact_audio=True)
    assert max(float(np.abs(r["raw"]).max()) for r in outs[0]) > 0.05
