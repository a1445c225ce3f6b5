"""Seeded scenarios of the hot path, driven through include/zl_b200.h on whichever library is given.

Each scenario returns a dict of numpy arrays; parity tests run it on the CUDA product and on the CPU
oracle and compare.  Shapes follow BASELINE.json configs / SURVEY.md §8d at sizes the oracle finishes
in seconds.
"""
import numpy as np

from libzl_b200 import _capi as K
from libzl_b200.engine import Engine


def noise_clip(seed, frames, nch=2):
    return np.random.default_rng(1000 + seed).uniform(-1, 1, (nch, frames)).astype(np.float32)


def sweep_clip(frames, rate=48000.0, f0=20.0, f1=20000.0):
    t = np.arange(frames) / rate
    T = frames / rate
    ph = 2 * np.pi * (f0 * t + (f1 - f0) * t * t / (2 * T))
    return np.stack([0.8 * np.sin(ph), 0.8 * np.sin(ph + np.pi / 2)]).astype(np.float32)


def collect(e, buses=True, master=True, levels=True, voices=True):
    out = {}
    S, B = e.cfg.n_sessions, e.cfg.buses_per_session
    if buses:
        for w, name in ((K.ZL_BUS_RAW, "raw"), (K.ZL_BUS_DRY, "dry"), (K.ZL_BUS_FX1, "fx1"), (K.ZL_BUS_FX2, "fx2")):
            if e.cfg.outputs & (K.ZL_OUT_RAW << w):
                out[name] = np.stack([np.stack(e.bus(s, b, w)) for s in range(S) for b in range(B)])
    if master and (e.cfg.outputs & K.ZL_OUT_MASTER):
        out["master"] = e.master_all()
    if levels:
        pk, sq = e.block_levels()
        out["peaks"], out["sumsq"] = pk, sq
    if voices:
        v = e.voices()
        out["v_position"] = np.array([x.position for x in v])
        out["v_env"] = np.array([x.envelope for x in v], np.float32)
        out["v_playing"] = np.array([x.playing for x in v])
        out["v_state"] = np.array([x.adsr_state for x in v])
        out["v_clip"] = np.array([x.clip for x in v])
        out["v_pos_gain"] = np.array([x.pos_gain for x in v], np.float32)
        out["v_pos_progress"] = np.array([x.pos_progress for x in v], np.float32)
    out["voice_frames"] = np.array([e.voice_frames()])
    return out


def append(acc, part):
    for k, v in part.items():
        acc.setdefault(k, []).append(v)


def finish(acc):
    res = {}
    for k, parts in acc.items():
        if k in ("raw", "dry", "fx1", "fx2", "master"):
            res[k] = np.concatenate(parts, axis=-1)
        elif k in ("peaks", "sumsq"):
            res[k] = np.concatenate(parts, axis=0)
        else:
            res[k] = np.stack(parts)
    return res


def cfg1_single_voice(lib, mode=K.ZL_MODE_Z, quirks=K.ZL_QUIRKS_FAITHFUL, looping=False, nblocks=24, nframes=512,
                      frames=10000, length_beats=-1.0):
    """configs[0]: one voice, one stereo 48 kHz clip, pitch 1.0, 512-frame blocks (test/playtest.py path)."""
    e = Engine(lib=lib, mode=mode, quirks=quirks)
    c = e.add_clip(noise_clip(1, frames))
    if length_beats != -1.0:
        e.set_clip_params(c, length_beats=length_beats)
    e.play(c, loop=looping)
    acc = {}
    done = 0
    while done < nblocks:
        nb = min(8, nblocks - done)
        e.render(nb, nframes, e.make_clocks(nb, nframes, first_block=done))
        append(acc, collect(e))
        done += nb
    e.close()
    return finish(acc)


def cfg2_polyphonic(lib, mode=K.ZL_MODE_Z, quirks=K.ZL_QUIRKS_FAITHFUL, n_sessions=1, buses=16, voices=8, nclips=8,
                    clip_frames=12000, nblocks=12, nframes=512, seed=2, launches=2, looping_fraction=0.4,
                    outputs=K.ZL_OUT_ALL):
    """configs[1] shape: buses x voices polyphony, random pitch 0.5-2.0, ADSR + pan, mixed to a stereo master."""
    rng = np.random.default_rng(seed)
    e = Engine(lib=lib, mode=mode, quirks=quirks, n_sessions=n_sessions, buses_per_session=buses, voices_per_bus=voices,
               max_clips=max(256, nclips * n_sessions), outputs=outputs)
    clips = []
    for s in range(n_sessions):
        row = []
        for k in range(nclips):
            pcm = noise_clip(seed * 100 + s * nclips + k, clip_frames) if k % 2 == 0 else sweep_clip(clip_frames)
            if k % 5 == 4:
                pcm = pcm[:1]          # a mono clip
            c = e.add_clip(pcm)
            e.set_clip_params(c, attack=float(rng.uniform(0, 0.05)), decay=float(rng.uniform(0, 0.2)),
                              sustain=float(rng.uniform(0.3, 1.0)), release=float(rng.uniform(0.01, 0.2)),
                              pan=float(rng.uniform(-1, 1)), volume_abs=float(rng.uniform(0.4, 1.0)),
                              start_s=float(rng.uniform(0, 0.05)), length_s=float(rng.uniform(0.1, 0.2)),
                              length_beats=float(rng.choice([1.5, 2.25, 0.75])), root_note=60)
            row.append(c)
        clips.append(row)
        for b in range(buses):
            e.set_passthrough(s, b, dry=float(rng.choice([1.0, 0.8])), fx1=float(rng.choice([1.0, 0.3, 0.0])),
                              fx2=float(rng.choice([0.0, 1.0])), pan=float(rng.choice([0.0, 0.0, -0.4, 0.6])),
                              muted=bool(rng.random() < 0.05))
            for v in range(voices):
                clip = row[int(rng.integers(nclips))]
                e.play(clip, loop=bool(rng.random() < looping_fraction), midi_channel=b - 2,
                       note=60 + int(rng.integers(-12, 13)), volume=float(rng.uniform(0.25, 1.0)),
                       slice_=(int(rng.integers(0, 16)) if rng.random() < 0.3 else None), session=s,
                       block=int(rng.integers(0, 3)))
    acc = {}
    for launch in range(launches):
        if launch == 1:   # stop a few voices and re-trigger others mid-way
            for s in range(n_sessions):
                for b in range(0, buses, 3):
                    e.stop(clips[s][b % nclips], midi_channel=b - 2, note=60 + b % 5, session=s, block=1)
                    e.play(clips[s][(b + 1) % nclips], midi_channel=b - 2, note=62, volume=0.6, session=s, block=2)
        e.render(nblocks, nframes, e.make_clocks(nblocks, nframes, first_block=launch * nblocks))
        append(acc, collect(e))
    e.close()
    return finish(acc)


def cfg3_sketch(lib, quirks=K.ZL_QUIRKS_FAITHFUL, nblocks=40, nframes=256, beat_aligned=True):
    """configs[2] shape: 10 track buses x looping clips with per-track meters; beat-aligned loops on a
    synthetic block clock (subbeat 1000 us so a 1-beat loop wraps every 18 blocks) and position loops."""
    rng = np.random.default_rng(3)
    e = Engine(lib=lib, quirks=quirks, buses_per_session=12)
    loop_frames = 6000
    clips = []
    for k in range(16):
        c = e.add_clip(noise_clip(300 + k, loop_frames + 500))
        lb = 1.0 if beat_aligned and k % 2 == 0 else 0.3   # 1 beat = 96 ticks x 1000 us = 96 ms = 4608 frames
        e.set_clip_params(c, length_s=loop_frames / 48000.0, length_beats=lb, pan=float(rng.uniform(-1, 1)),
                          volume_abs=0.8, attack=0.0, decay=0.1, sustain=1.0, release=0.05)
        clips.append(c)
    for t in range(10):
        e.set_passthrough(0, t + 2, dry=0.8, fx1=0.3, fx2=0.0, pan=float(rng.uniform(-0.5, 0.5)))
        for k in range(8):
            e.play(clips[(t + k) % 16], loop=True, midi_channel=t, note=60 + (k % 3) * 12 - 12, volume=0.5, tick=96 * 4)
    acc = {}
    period = int(round(nframes * 1e6 / 48000.0))
    done = 0
    ticks = []
    while done < nblocks:
        nb = min(10, nblocks - done)
        clocks = e.make_clocks(nb, nframes, first_block=done, playhead=96 * 4, playhead_usecs=1_000_000, subbeat_usecs=1000)
        if done == 20:   # queue the other clips at a "bar boundary"; stop some
            for t in range(0, 10, 2):
                e.stop(clips[t % 16], midi_channel=t, note=48, block=0)
                e.play(clips[(t + 9) % 16], loop=True, midi_channel=t, note=55, volume=0.4, tick=96 * 4 + (done * period) // 1000, block=1)
        e.render(nb, nframes, clocks)
        part = collect(e)
        lv = []
        for t in range(10):
            x = e.levels_tick(0, t + 2, nb - 1)
            lv.append([x.peak_a, x.peak_b])
        part["meter_peaks"] = np.array(lv)
        gp = [e.positions(c) for c in clips[:4]]
        part["pos_counts"] = np.array([len(g[0]) for g in gp])
        append(acc, part)
        done += nb
    e.close()
    return finish(acc)


def compare(a, b, audio_tol=1e-5, exact_audio=False, loose_levels=False):
    """Returns a list of mismatch strings (empty = parity)."""
    bad = []
    for k in a:
        x, y = np.asarray(a[k]), np.asarray(b[k])
        if x.shape != y.shape:
            bad.append(f"{k}: shape {x.shape} vs {y.shape}")
            continue
        if k in ("raw", "dry", "fx1", "fx2", "master", "v_env", "v_pos_gain", "v_pos_progress"):
            if exact_audio:
                if not np.array_equal(x, y):
                    bad.append(f"{k}: not bit-exact, max abs {np.abs(x - y).max():.3g}")
            else:
                err = float(np.abs(x.astype(np.float64) - y).max()) if x.size else 0.0
                if not err <= audio_tol:
                    bad.append(f"{k}: max abs err {err:.3g} > {audio_tol}")
        elif loose_levels and k in ("peaks", "meter_peaks"):
            # two different roundings of the same audio (FMA-contracted vs not): 131072 * 1e-5 ~ 1.3 counts
            if np.abs(x.astype(np.int64) - y).max() > 4:
                bad.append(f"{k}: peaks differ by {np.abs(x.astype(np.int64) - y).max()}")
        elif k == "sumsq":
            if not np.allclose(x, y, rtol=1e-3 if loose_levels else 1e-9, atol=1e-12):
                bad.append(f"{k}: rel err {np.abs(x - y).max():.3g}")
        else:
            if not np.array_equal(x, y):
                bad.append(f"{k}: integer/state mismatch ({int((x != y).sum())} of {x.size})")
    return bad


def adsr_via_voice(lib, sr, a, d, s, r, n_on, n_total, renote=False, period=100):
    """A juce::ADSR sequence (noteOn at 0, noteOff at frame n_on; renote: noteOff again before every later frame) taken
    through the VOICE path of `lib`: a DC clip of 1.0 at pitch ratio 1, velocity 1, so the bus carries the envelope itself.
    Plain sequences run in J-mode (stop command at the period that starts at n_on); renote sequences run in Z-mode with
    only quirk Q3 on, where the one-shot's release window (lib/SamplerSynthVoice.cpp:253-256) starts after frame n_on - 1
    and the voice ends at the stop position.  Returns (envelope, frames that are comparable)."""
    a, d, s, r = (float(np.float32(x)) for x in (a, d, s, r))
    n_on, n_total = int(n_on), int(n_total)
    assert n_on % period == 0 and n_total % period == 0
    nb = n_total // period
    if not renote:
        e = Engine(lib=lib, mode=K.ZL_MODE_J, sample_rate=float(sr), max_frames=period, max_blocks=nb)
        c = e.add_clip(np.ones((2, n_total + 16), np.float32), rate=float(sr))
        e.set_clip_params(c, attack=a, decay=d, sustain=s, release=r, root_note=60)
        e.play(c, midi_channel=-2, note=60, volume=1.0)
        e.stop(c, midi_channel=-2, note=60, block=n_on // period)
        e.render(nb, period)
        l, rr = e.bus(0, 0, K.ZL_BUS_RAW)
        e.close()
        assert np.array_equal(l, rr)
        return l, n_total
    rel_frames = int(np.ceil(np.float64(np.float32(r)) * sr))
    stop = n_on + rel_frames - 1                     # relStart = stop - release * rate lies in (n_on - 1, n_on]
    assert stop - np.float64(np.float32(r)) * sr <= n_on and stop - np.float64(np.float32(r)) * sr > n_on - 1
    e = Engine(lib=lib, mode=K.ZL_MODE_Z, quirks=K.ZL_Q3_RENOTEOFF, sample_rate=float(sr), max_frames=period, max_blocks=nb)
    c = e.add_clip(np.ones((2, n_total + 16), np.float32), rate=float(sr))
    e.set_clip_params(c, attack=a, decay=d, sustain=s, release=r, root_note=60, pan=0.0, volume_abs=1.0,
                      start_s=0.0, length_s=(stop + 0.5) / sr)
    assert e.slice_bounds(c, -1) == (0, stop)
    e.play(c, midi_channel=-2, note=60, volume=1.0)
    e.render(nb, period)
    l, rr = e.bus(0, 0, K.ZL_BUS_RAW)
    e.close()
    assert np.array_equal(l, rr)
    return l * np.float32(2.0), min(stop, n_total)    # M/S pan at pan 0: l = 0.5 * m + 0 (exact halving)
