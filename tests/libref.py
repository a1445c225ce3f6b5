"""The reference's OWN compiled hot path (lib/SamplerSynthVoice.cpp, lib/SamplerSynth.cpp, lib/JackPassthrough.cpp,
lib/ClipAudioSourcePositionsModel.cpp, built in place into oracle/_ref/libzl_juceref.so with the stand-ins of
oracle/ref/shim) driven from Python, and the same operation list replayed through include/zl_b200.h on the oracle or the
product.  TEST INFRASTRUCTURE.

The reference objects are process-wide singletons (SamplerSynth::instance(), 12 SamplerChannels x 8 voices), so every
scenario runs in a fresh subprocess.

An operation list is a list of tuples:
  ("clip", pcm[nch, n] float32, rate, {zl_clip_params fields})      clips get ids 0, 1, ... in order
  ("passthrough", bus, dry, fx1, fx2, pan, muted)
  ("enable", midi_channel, enabled)
  ("cmd", block_offset, {zl_clip_command fields}, tick)               applied at the start of that period of the NEXT process
  ("process", nblocks, nframes, {first_block, playhead, playhead_usecs, subbeat_usecs, t0_usecs, rate})
  ("meters", [(audiolevels_channel, bus), ...])                       AudioLevels channel (1 = system playback, 3.. = tracks) fed
                                                                      by the dry output of that bus; one meter tick follows every
                                                                      process op (AudioLevels::timerCallback, 20 Hz in the reference)
Results: one dict per process op with raw / dry / fx1 / fx2 as [12, 2, nblocks * nframes] float32, and per clip the
positions model after the op: peak_gain (ClipAudioSourcePositionsModel::peakGain, hysteresis included) and the sorted
(gain, progress) pairs of the live positions.
"""
import ctypes as C
import multiprocessing as mp
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUSES, VOICES = 12, 8
f32p = C.POINTER(C.c_float)


def clocks(nblocks, nframes, first_block=0, playhead=0, playhead_usecs=None, subbeat_usecs=5208, t0_usecs=1_000_000, rate=48000.0):
    period = int(round(nframes * 1e6 / rate))
    out = []
    for i in range(nblocks):
        cur = t0_usecs + (first_block + i) * period
        out.append((cur, cur + period, playhead, t0_usecs if playhead_usecs is None else playhead_usecs, subbeat_usecs))
    return out


# ---------------------------------------------------------------------------------------------------------------
def _ref_worker(ops, sample_rate, max_frames, q):
    try:
        sys.path.insert(0, ROOT)
        from oracle import loader
        d = C.CDLL(loader.JUCEREF_SO)
        d.zlref_init.argtypes = [C.c_double, C.c_int]
        d.zlref_clip_new.argtypes = [C.POINTER(f32p), C.c_int, C.c_longlong, C.c_double]
        d.zlref_clip_set.argtypes = [C.c_int] + [C.c_float] * 9 + [C.c_int, C.POINTER(C.c_double), C.c_int]
        d.zlref_command.argtypes = [C.c_int] * 9 + [C.c_int, C.c_float, C.c_ulonglong]
        d.zlref_channel_enabled.argtypes = [C.c_int, C.c_int]
        d.zlref_passthrough_set.argtypes = [C.c_int] + [C.c_float] * 4 + [C.c_int]
        d.zlref_process.argtypes = [C.c_int] + [C.c_ulonglong] * 5
        d.zlref_read.argtypes = [C.c_int, C.c_int, f32p, f32p, C.c_int]
        d.zlref_clip_peak_gain.restype, d.zlref_clip_peak_gain.argtypes = C.c_float, [C.c_int]
        d.zlref_clip_positions.argtypes = [C.c_int, C.POINTER(C.c_longlong), f32p, f32p]
        d.zlref_connect.argtypes = [C.c_char_p, C.c_char_p]
        d.zlref_levels_tick.argtypes = [C.POINTER(C.c_int), f32p]
        assert d.zlref_init(float(sample_rate), int(max_frames)) == 0
        d.zlref_sched_schedule.argtypes = [C.c_int] * 10 + [C.c_float, C.c_ulonglong]
        d.zlref_sched_channel_enabled.argtypes = [C.c_ulonglong, C.c_int, C.c_int]
        d.zlref_sched_cycle.argtypes = [C.c_int, C.c_ulonglong, C.c_ulonglong, C.POINTER(C.c_ulonglong), C.POINTER(C.c_int), f32p, C.c_int]
        d.zlref_sched_state.argtypes = [C.POINTER(C.c_ulonglong)]
        nclips, pending, results = 0, [], []
        sched_period = 0
        meters = []
        lv_names = ["AudioLevels-SystemCapture", "AudioLevels-SystemPlayback", "AudioLevels-SystemRecorder"] + [f"AudioLevels-Channel{i}" for i in range(1, 11)]
        clip_par = {}

        def set_params(cid, par):
            clip_par[cid].update(par)
            p = clip_par[cid]
            sl = p.get("slice_pos")
            if sl is None:
                n_sl = int(p.get("n_slices", 16))
                sl = [0.0] + [(1.0 / n_sl) * (k + 1) for k in range(n_sl - 1)] if n_sl > 0 else []
                # lib/ClipAudioSource.cpp:495-528 from an empty list: 0, then pos += inc starting at inc
                sl, pos, inc = [0.0], 1.0 / n_sl, 1.0 / n_sl
                while len(sl) < n_sl:
                    sl.append(pos); pos += inc
            arr = (C.c_double * len(sl))(*sl)
            d.zlref_clip_set(cid, p["start_s"], p["length_s"], p["length_beats"], p["pan"], p["volume_abs"], p["attack"], p["decay"],
                             p["sustain"], p["release"], int(p["root_note"]), arr, len(sl))

        for op in ops:
            if op[0] == "clip":
                _, pcm, rate, par = op
                pcm = np.ascontiguousarray(np.atleast_2d(pcm), np.float32)
                ptrs = (f32p * pcm.shape[0])(*[pcm[i].ctypes.data_as(f32p) for i in range(pcm.shape[0])])
                cid = d.zlref_clip_new(ptrs, pcm.shape[0], pcm.shape[1], float(rate))
                assert cid == nclips
                nclips += 1
                dur = np.float32(np.float64(pcm.shape[1]) / rate)
                clip_par[cid] = dict(start_s=0.0, length_s=float(dur), length_beats=-1.0, pan=0.0,
                                     volume_abs=float(np.exp(np.float32(-9.0) * np.float32(1.0 / 20.0), dtype=np.float32)),
                                     attack=0.0, decay=0.1, sustain=1.0, release=0.05, root_note=60, n_slices=16)
                set_params(cid, par)
            elif op[0] == "passthrough":
                d.zlref_passthrough_set(*[op[1]] + [float(x) for x in op[2:6]] + [int(op[6])])
            elif op[0] == "enable":
                d.zlref_channel_enabled(int(op[1]), int(op[2]))
            elif op[0] == "meters":
                assert d.zlref_levels_init() == 0
                for chan, bus in op[1]:
                    for side, port in (("left_in", "dryOutLeft"), ("right_in", "dryOutRight")):
                        assert d.zlref_connect(f"ZLRefPassthrough{bus}:{port}".encode(), f"{lv_names[chan]}:{side}".encode()) == 0
                    meters.append((chan, bus))
            elif op[0] == "cmd":
                pending.append(op[1:])
            elif op[0] == "sched_start":
                d.zlref_sched_start(int(op[1]))
            elif op[0] == "sched_stop":
                d.zlref_sched_stop()
            elif op[0] == "sched_bpm":
                d.zlref_sched_set_bpm(int(op[1]))
            elif op[0] == "sched_cmd":
                g = dict(clip=-1, midi_note=-1, midi_channel=-1, start_playback=0, stop_playback=0, change_slice=0, slice=-1,
                         change_looping=0, looping=0, change_volume=0, volume=0.0)
                g.update(op[2])
                assert d.zlref_sched_schedule(g["clip"], g["midi_note"], g["midi_channel"], g["start_playback"], g["stop_playback"], g["change_slice"],
                                              g["slice"], g["change_looping"], g["looping"], g["change_volume"], float(g["volume"]), int(op[1])) == 0
            elif op[0] == "sched_queue_start":
                d.zlref_sched_queue_start(int(op[1]), int(op[2]))
            elif op[0] == "sched_queue_stop":
                d.zlref_sched_queue_stop(int(op[1]), int(op[2]))
            elif op[0] == "sched_enable":
                d.zlref_sched_channel_enabled(int(op[1]), int(op[2]), int(op[3]))
            elif op[0] == "sched_run":
                _, nper, nf = op
                n = nper * nf
                out = {k: np.zeros((BUSES, 2, n), np.float32) for k in ("raw", "dry", "fx1", "fx2")}
                out["clocks"], out["commands"] = [], []
                period = int(round(nf * 1e6 / sample_rate))
                l, r = np.zeros(nf, np.float32), np.zeros(nf, np.float32)
                for b in range(nper):
                    cur = 1_000_000 + sched_period * period
                    sched_period += 1
                    o3 = (C.c_ulonglong * 3)(); ci = (C.c_int * (11 * 256))(); cv = (C.c_float * 256)()
                    m = d.zlref_sched_cycle(nf, cur, cur + period, o3, ci, cv, 256)
                    assert 0 <= m <= 256
                    out["clocks"].append((cur, cur + period, int(o3[0]), int(o3[1]), int(o3[2])))
                    out["commands"].append([tuple(ci[11 * k:11 * k + 10]) + (float(cv[k]),) for k in range(m)])
                    for w, k in enumerate(("raw", "dry", "fx1", "fx2")):
                        for bus in range(BUSES):
                            d.zlref_read(bus, w, l.ctypes.data_as(f32p), r.ctypes.data_as(f32p), nf)
                            out[k][bus, 0, b * nf:(b + 1) * nf] = l
                            out[k][bus, 1, b * nf:(b + 1) * nf] = r
                st = (C.c_ulonglong * 4)()
                d.zlref_sched_state(st)
                out["state"] = tuple(int(x) for x in st)
                out["positions"] = []
                results.append(out)
            elif op[0] == "process":
                _, nb, nf, ck = op
                n = nb * nf
                out = {k: np.zeros((BUSES, 2, n), np.float32) for k in ("raw", "dry", "fx1", "fx2")}
                for b, (cur, nxt, ph, phu, sbu) in enumerate(clocks(nb, nf, **ck)):
                    for blk, f, tick in [c for c in pending if c[0] == b]:
                        g = dict(clip=-1, midi_note=-1, midi_channel=-1, start_playback=0, stop_playback=0, change_slice=0, slice=-1,
                                 change_looping=0, looping=0, change_volume=0, volume=0.0)
                        g.update(f)
                        d.zlref_command(g["clip"], g["midi_note"], g["midi_channel"], g["start_playback"], g["stop_playback"], g["change_slice"],
                                        g["slice"], g["change_looping"], g["looping"], g["change_volume"], float(g["volume"]), int(tick))
                    assert d.zlref_process(nf, cur, nxt, ph, phu, sbu) == 0
                    l, r = np.zeros(nf, np.float32), np.zeros(nf, np.float32)
                    for w, k in enumerate(("raw", "dry", "fx1", "fx2")):
                        for bus in range(BUSES):
                            d.zlref_read(bus, w, l.ctypes.data_as(f32p), r.ctypes.data_as(f32p), nf)
                            out[k][bus, 0, b * nf:(b + 1) * nf] = l
                            out[k][bus, 1, b * nf:(b + 1) * nf] = r
                pending = [c for c in pending if c[0] >= nb]          # late commands: dropped by the boundary, never sent here
                pending = []
                pos = []
                for cid in range(nclips):
                    ids = (C.c_longlong * 32)(); g = np.zeros(32, np.float32); p = np.zeros(32, np.float32)
                    m = d.zlref_clip_positions(cid, ids, g.ctypes.data_as(f32p), p.ctypes.data_as(f32p))
                    live = sorted((float(g[i]), float(p[i])) for i in range(m) if ids[i] > -1)
                    pos.append((float(d.zlref_clip_peak_gain(cid)), live))
                out["positions"] = pos
                if meters:
                    pk = (C.c_int * 26)(); lv = np.zeros((13, 6), np.float32)
                    d.zlref_levels_tick(pk, lv.ctypes.data_as(f32p))
                    out["meters"] = [(pk[2 * c], pk[2 * c + 1]) + tuple(float(x) for x in lv[c, :5]) for c, _ in meters]
                results.append(out)
        q.put(("ok", results))
    except Exception as ex:      # noqa: BLE001
        import traceback
        q.put(("error", traceback.format_exc()))


def run_reference(ops, sample_rate=48000.0, max_frames=1024):
    """Runs the operation list on the reference-compiled objects in a fresh process."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    p = ctx.Process(target=_ref_worker, args=(ops, sample_rate, max_frames, q))
    p.start()
    import queue
    status, payload = "error", "the reference process died without an answer"
    for _ in range(1200):
        try:
            status, payload = q.get(timeout=0.5)
            break
        except queue.Empty:
            if not p.is_alive():
                break
    p.join(timeout=10)
    if p.is_alive():
        p.terminate()
    if status != "ok":
        raise RuntimeError("reference run failed:\n" + payload)
    return payload


def available():
    sys.path.insert(0, ROOT)
    from oracle import loader
    if not os.path.exists(loader.JUCEREF_SO):
        return False
    try:
        return hasattr(C.CDLL(loader.JUCEREF_SO), "zlref_init")
    except OSError:
        return False


# ---------------------------------------------------------------------------------------------------------------
def run_engine(lib, ops, sample_rate=48000.0, max_frames=1024, max_blocks=64, sched_batch=1):
    """The same operation list through include/zl_b200.h (oracle or product): one session of 12 buses x 8 voices,
    Z-mode, every quirk faithful."""
    from libzl_b200 import _capi as K
    from libzl_b200.engine import Engine
    e = Engine(lib=lib, n_sessions=1, buses_per_session=BUSES, voices_per_bus=VOICES, mode=K.ZL_MODE_Z, quirks=K.ZL_QUIRKS_FAITHFUL,
               outputs=K.ZL_OUT_ALL, sample_rate=float(sample_rate), max_frames=max_frames, max_blocks=max_blocks)
    results, nclips, meters = [], 0, []
    sched, sched_period = None, 0
    if any(op[0].startswith("sched_") for op in ops):
        sched = C.c_void_p()
        assert lib.sched_create(float(sample_rate), int(max_frames), C.byref(sched)) == K.ZL_OK
    for op in ops:
        if op[0] == "meters":
            meters = list(op[1])
        elif op[0] == "sched_start":
            assert lib.sched_start(sched, int(op[1])) == K.ZL_OK
        elif op[0] == "sched_stop":
            assert lib.sched_stop(sched) == K.ZL_OK
        elif op[0] == "sched_bpm":
            assert lib.sched_set_bpm(sched, int(op[1])) == K.ZL_OK
        elif op[0] == "sched_cmd":
            c = K.zl_clip_command()
            lib.clip_command_default(C.byref(c))
            for k_, v_ in op[2].items():
                setattr(c, k_, v_)
            assert lib.sched_schedule(sched, C.byref(c), int(op[1])) == K.ZL_OK
        elif op[0] == "sched_queue_start":
            assert lib.sched_queue_start(sched, int(op[1]), int(op[2])) == K.ZL_OK
        elif op[0] == "sched_queue_stop":
            assert lib.sched_queue_stop(sched, int(op[1]), int(op[2])) == K.ZL_OK
        elif op[0] == "sched_enable":
            assert lib.sched_schedule_channel_enabled(sched, int(op[1]), int(op[2]), int(op[3])) == K.ZL_OK
        elif op[0] == "sched_run":
            _, nper, nf = op
            period = int(round(nf * 1e6 / sample_rate))
            out = {k_: [] for k_ in ("raw", "dry", "fx1", "fx2")}
            out["clocks"], out["commands"] = [], []
            done = 0
            while done < nper:
                nb = min(sched_batch, nper - done, max_blocks)
                ck = (K.zl_block_clock * nb)()
                for i in range(nb):
                    ck[i].current_usecs = 1_000_000 + (sched_period + i) * period
                    ck[i].next_usecs = ck[i].current_usecs + period
                ev = (K.zl_sched_event * 4096)()
                cnt = C.c_int32()
                assert lib.sched_advance(sched, nb, nf, ck, ev, 4096, C.byref(cnt)) == K.ZL_OK
                per_block = [[] for _ in range(nb)]
                events = [ev[i] for i in range(cnt.value)]
                # a channel switched on or off takes effect from its own period on: render the batch in segments that start there
                cuts = sorted({0, nb} | {x.block for x in events if x.kind != K.ZL_SCHED_CLIP_COMMAND})
                for i in range(nb):
                    out["clocks"].append((ck[i].current_usecs, ck[i].next_usecs, ck[i].playhead, ck[i].playhead_usecs, ck[i].subbeat_usecs))
                for lo, hi in zip(cuts[:-1], cuts[1:]):
                    for x in events:
                        if not (lo <= x.block < hi):
                            continue
                        if x.kind == K.ZL_SCHED_CLIP_COMMAND:
                            c = x.cmd
                            per_block[x.block].append((c.clip, c.midi_note, c.midi_channel, c.start_playback, c.stop_playback, c.change_slice, c.slice,
                                                       c.change_looping, c.looping, c.change_volume, float(c.volume)))
                            e._chk(lib.command(e.ctx, 0, x.block - lo, x.tick, C.byref(c)), "command")
                        else:
                            assert x.block == lo
                            e.set_bus_enabled(0, x.channel + 2, bool(x.enabled))
                    seg = (K.zl_block_clock * (hi - lo))(*[ck[i] for i in range(lo, hi)])
                    e.render(hi - lo, nf, seg)
                    for w, k_ in ((K.ZL_BUS_RAW, "raw"), (K.ZL_BUS_DRY, "dry"), (K.ZL_BUS_FX1, "fx1"), (K.ZL_BUS_FX2, "fx2")):
                        out[k_].append(np.stack([np.stack(e.bus(0, b, w)) for b in range(BUSES)]))
                out["commands"] += per_block
                sched_period += nb
                done += nb
            for k_ in ("raw", "dry", "fx1", "fx2"):
                out[k_] = np.concatenate(out[k_], axis=-1)
            st = [C.c_uint64() for _ in range(3)] + [C.c_uint64(), C.c_int32(), C.c_uint64()]
            lib.sched_state(sched, C.byref(st[0]), C.byref(st[1]), C.byref(st[2]), C.byref(st[3]), C.byref(st[4]), C.byref(st[5]))
            out["state"] = (int(st[3].value), int(st[4].value), int(st[5].value))
            out["positions"] = []
            results.append(out)
        elif op[0] == "clip":
            _, pcm, rate, par = op
            cid = e.add_clip(pcm, rate)
            assert cid == nclips
            nclips += 1
            par = dict(par)
            sl = par.pop("slice_pos", None)
            if par or sl is not None:
                e.set_clip_params(cid, slice_pos=sl, **par)
        elif op[0] == "passthrough":
            e.set_passthrough(0, op[1], dry=op[2], fx1=op[3], fx2=op[4], pan=op[5], muted=bool(op[6]))
        elif op[0] == "enable":
            e.set_bus_enabled(0, op[1] + 2, bool(op[2]))
        elif op[0] == "cmd":
            _, blk, f, tick = op
            e.command(0, blk, tick, **f)
        elif op[0] == "process":
            _, nb, nf, ck = op
            arr = (K.zl_block_clock * nb)()
            for i, (cur, nxt, ph, phu, sbu) in enumerate(clocks(nb, nf, **ck)):
                arr[i].current_usecs, arr[i].next_usecs, arr[i].playhead, arr[i].playhead_usecs, arr[i].subbeat_usecs = cur, nxt, ph, phu, sbu
            e.render(nb, nf, arr)
            out = {}
            for w, k in ((K.ZL_BUS_RAW, "raw"), (K.ZL_BUS_DRY, "dry"), (K.ZL_BUS_FX1, "fx1"), (K.ZL_BUS_FX2, "fx2")):
                out[k] = np.stack([np.stack(e.bus(0, b, w)) for b in range(BUSES)])
            pos = []
            for cid in range(nclips):
                g, p = e.positions(cid)
                peak = e.clip_peak_gain(cid) if hasattr(e, "clip_peak_gain") else None
                pos.append((peak, sorted((float(a), float(b)) for a, b in zip(g, p))))
            out["positions"] = pos
            if meters:
                out["meters"] = []
                for chan, bus in meters:
                    x = e.levels_tick(0, bus, nb - 1, is_playback=(chan == 1))
                    out["meters"].append((x.peak_a, x.peak_b, x.db_a, x.db_b, x.db_sum, x.hold_a, x.hold_b))
            results.append(out)
    e.close()
    if sched is not None:
        lib.sched_destroy(sched)
    return results


def disabled_buses(ops):
    """Per process op, the buses switched off at that time.  The reference leaves the JACK buffers of a disabled channel
    untouched (whatever the server's memory holds, lib/SamplerSynth.h:47-53); the boundary defines them as silence - the one
    place the two are not comparable."""
    off, out = set(), []
    for op in ops:
        if op[0] == "enable":
            (off.discard if op[2] else off.add)(op[1] + 2)
        elif op[0] == "sched_enable":                 # takes effect some steps into the next run: that run is skipped for the bus
            off.add(op[2] + 2)
            if op[3]:
                pending_on = op[2] + 2
                out_mark = ("on", pending_on)
                off.add(pending_on)
        elif op[0] in ("process", "sched_run"):
            out.append(sorted(off))
            for b in [x for x in off if any(o[0] == "sched_enable" and o[2] + 2 == x and o[3] for o in ops[:ops.index(op)])]:
                # re-enabled by a timer command during this run: comparable again from the next run on
                last = [o for o in ops[:ops.index(op)] if o[0] == "sched_enable" and o[2] + 2 == b][-1]
                if last[3]:
                    off.discard(b)
    return out


def compare(a, b, check_positions=True, skip_buses=None):
    """Mismatch strings between two result lists (empty = parity, bit for bit)."""
    bad = []
    if len(a) != len(b):
        return [f"{len(a)} vs {len(b)} process ops"]
    for i, (x, y) in enumerate(zip(a, b)):
        keep = np.ones(BUSES, bool)
        if skip_buses:
            keep[skip_buses[i]] = False
        for k in ("raw", "dry", "fx1", "fx2"):
            if x[k].shape != y[k].shape:
                bad.append(f"op {i} {k}: shape")
            elif not np.array_equal(x[k][keep], y[k][keep]):
                bad.append(f"op {i} {k}: not bit-exact, max abs {np.abs(x[k] - y[k]).max():.3g}, first at {np.argwhere(x[k] != y[k])[0].tolist()}")
        if "clocks" in x or "clocks" in y:
            if x["clocks"] != y["clocks"]:
                k_ = next(j for j, (a_, b_) in enumerate(zip(x["clocks"], y["clocks"])) if a_ != b_)
                bad.append(f"op {i}: block clocks differ first at period {k_}: {x['clocks'][k_]} vs {y['clocks'][k_]}")
            if x["commands"] != y["commands"]:
                k_ = next(j for j, (a_, b_) in enumerate(zip(x["commands"], y["commands"])) if a_ != b_)
                bad.append(f"op {i}: command stream differs first at period {k_}: {x['commands'][k_]} vs {y['commands'][k_]}")
            if tuple(x["state"][:3]) != tuple(y["state"][:3]):
                bad.append(f"op {i}: timer state (cumulativeBeat, beat, scheduleAhead) {x['state']} vs {y['state']}")
        if "meters" in x or "meters" in y:
            for c, (ma, mb) in enumerate(zip(x.get("meters", []), y.get("meters", []))):
                if ma[:2] != mb[:2]:
                    bad.append(f"op {i} meter {c}: integer peaks {ma[:2]} vs {mb[:2]}")
                if not np.allclose(ma[2:], mb[2:], rtol=1e-6, atol=1e-4):          # dB through libm log10f / pow
                    bad.append(f"op {i} meter {c}: levels {ma[2:]} vs {mb[2:]}")
        if check_positions:
            for c, ((pa, la), (pb, lb)) in enumerate(zip(x["positions"], y["positions"])):
                if la != lb:
                    bad.append(f"op {i} clip {c}: positions {la[:3]} vs {lb[:3]}")
                if pa is not None and pb is not None and pa != pb:
                    bad.append(f"op {i} clip {c}: peakGain {pa} vs {pb}")
    return bad


# ---------------------------------------------------------------------------------------------------------------
# seeded scenarios (shapes of BASELINE configs[0..2] inside the reference's fixed 12 x 8 layout)
def scenario_single_voice(looping=False, nblocks=24, nframes=512, frames=10000, length_beats=-1.0):
    pcm = np.random.default_rng(1001).uniform(-1, 1, (2, frames)).astype(np.float32)
    par = {} if length_beats == -1.0 else {"length_beats": length_beats}
    ops = [("clip", pcm, 48000.0, par),
           ("cmd", 0, dict(clip=0, midi_channel=-2, midi_note=60, change_volume=1, volume=1.0, looping=int(looping),
                           stop_playback=int(looping), start_playback=1), 0)]
    done = 0
    while done < nblocks:
        nb = min(8, nblocks - done)
        ops.append(("process", nb, nframes, dict(first_block=done)))
        done += nb
    return ops


def scenario_polyphonic(seed=2, nclips=8, clip_frames=12000, nblocks=12, nframes=512, launches=4, looping_fraction=0.4):
    rng = np.random.default_rng(seed)
    ops = []
    for k in range(nclips):
        if k % 2 == 0:
            pcm = np.random.default_rng(1000 + seed * 100 + k).uniform(-1, 1, (2, clip_frames)).astype(np.float32)
        else:
            t = np.arange(clip_frames) / 48000.0
            ph = 2 * np.pi * (20.0 * t + (20000.0 - 20.0) * t * t / (2 * clip_frames / 48000.0))
            pcm = np.stack([0.8 * np.sin(ph), 0.8 * np.sin(ph + np.pi / 2)]).astype(np.float32)
        if k % 5 == 4:
            pcm = pcm[:1]
        ops.append(("clip", pcm, 48000.0 if k % 3 else 44100.0, dict(
            attack=float(rng.uniform(0, 0.05)), decay=float(rng.uniform(0, 0.2)), sustain=float(rng.uniform(0.3, 1.0)),
            release=float(rng.uniform(0.01, 0.2)), pan=float(rng.uniform(-1, 1)), volume_abs=float(rng.uniform(0.4, 1.0)),
            start_s=float(rng.uniform(0, 0.05)), length_s=float(rng.uniform(0.1, 0.2)),
            length_beats=float(rng.choice([1.5, 2.25, 0.75])), root_note=60)))
    for b in range(BUSES):
        ops.append(("passthrough", b, float(rng.choice([1.0, 0.8])), float(rng.choice([1.0, 0.3, 0.0])), float(rng.choice([0.0, 1.0])),
                    float(rng.choice([0.0, 0.0, -0.4, 0.6])), int(rng.random() < 0.08)))
        for v in range(VOICES + 1):                                      # one more than fits: the ninth start is dropped
            loop = bool(rng.random() < looping_fraction)
            f = dict(clip=int(rng.integers(nclips)), midi_channel=b - 2, midi_note=60 + int(rng.integers(-12, 13)), change_volume=1,
                     volume=float(rng.uniform(0.25, 1.0)), looping=int(loop), stop_playback=int(loop), start_playback=1)
            if rng.random() < 0.3:
                f.update(change_slice=1, slice=int(rng.integers(0, 16)))
            ops.append(("cmd", int(rng.integers(0, 3)), f, 0))
    for launch in range(launches):
        if launch >= 1:      # stops, restarts of playing notes (merge path), volume updates, fresh one-shots
            for b in range(0, BUSES, 2):
                ops.append(("cmd", 1, dict(clip=int(rng.integers(nclips)), midi_channel=b - 2, midi_note=60 + int(rng.integers(-12, 13)), stop_playback=1), 0))
                ops.append(("cmd", 2, dict(clip=(b + launch) % nclips, midi_channel=b - 2, midi_note=62, change_volume=1, volume=0.6, start_playback=1), 0))
                ops.append(("cmd", 3, dict(clip=(b + launch) % nclips, midi_channel=b - 2, midi_note=62, change_volume=1, volume=0.35), 0))
        if launch == 2:
            ops.append(("enable", 3, 0))      # SamplerSynth::setChannelEnabled: commands still drain, voices stand still
        if launch == 3:
            ops.append(("enable", 3, 1))
        ops.append(("process", nblocks, nframes, dict(first_block=launch * nblocks)))
    return ops


def scenario_sketch(beat_aligned=True, nblocks=40, nframes=256):
    rng = np.random.default_rng(3)
    ops, loop_frames = [], 6000
    for k in range(16):
        pcm = np.random.default_rng(1300 + k).uniform(-1, 1, (2, loop_frames + 500)).astype(np.float32)
        lb = 1.0 if beat_aligned and k % 2 == 0 else 0.3
        ops.append(("clip", pcm, 48000.0, dict(length_s=loop_frames / 48000.0, length_beats=lb, pan=float(rng.uniform(-1, 1)),
                                                volume_abs=0.8, attack=0.0, decay=0.1, sustain=1.0, release=0.05)))
    ops.append(("meters", [(1, 0)] + [(3 + t, t + 2) for t in range(10)]))          # playback meter on the global bus, ten track meters
    ops.append(("cmd", 2, dict(clip=0, midi_channel=-2, midi_note=67, change_volume=1, volume=0.9, start_playback=1), 0))
    for t in range(10):
        ops.append(("passthrough", t + 2, 0.8, 0.3, 0.0, float(rng.uniform(-0.5, 0.5)), 0))
        for k in range(8):
            ops.append(("cmd", 0, dict(clip=(t + k) % 16, midi_channel=t, midi_note=60 + (k % 3) * 12 - 12, change_volume=1, volume=0.5,
                                       looping=1, stop_playback=1, start_playback=1), 96 * 4))
    period = int(round(nframes * 1e6 / 48000.0))
    done = 0
    while done < nblocks:
        nb = min(10, nblocks - done)
        if done == 20:
            for t in range(0, 10, 2):
                ops.append(("cmd", 0, dict(clip=t % 16, midi_channel=t, midi_note=48, stop_playback=1), 0))
                ops.append(("cmd", 1, dict(clip=(t + 9) % 16, midi_channel=t, midi_note=55, change_volume=1, volume=0.4, looping=1,
                                           stop_playback=1, start_playback=1), 96 * 4 + (done * period) // 1000))
        ops.append(("process", nb, nframes, dict(first_block=done, playhead=96 * 4, playhead_usecs=1_000_000, subbeat_usecs=1000)))
        done += nb
    return ops


def scenario_scheduler(nframes=256):
    """Row n2: a sketch driven by the command scheduler instead of a hand-written command stream.  Commands scheduled some
    steps ahead while the timer is paused and while it runs, two equivalent commands landing on one step (merged), clips queued
    to start on the next bar, a queued stop that removes what is still waiting, a tempo change, a channel switched off by a
    timer command, stop() with commands still in the ring; beat-aligned loops, so a wrong tick is audible."""
    ops = []
    for k in range(4):
        pcm = np.random.default_rng(1400 + k).uniform(-1, 1, (2, 30000)).astype(np.float32)
        ops.append(("clip", pcm, 48000.0, dict(length_s=0.5, length_beats=1.0 if k % 2 == 0 else 0.75, pan=0.2 * k - 0.3, volume_abs=0.8,
                                                attack=0.001, decay=0.05, sustain=0.8, release=0.02)))
    play = lambda clip, ch, note=60, vol=0.7, loop=1: dict(clip=clip, midi_channel=ch, midi_note=note, change_volume=1, volume=vol, looping=loop,
                                                           stop_playback=loop, start_playback=1)
    # paused: delays count from the ring's read head
    ops += [("sched_cmd", 0, play(0, -2)), ("sched_cmd", 3, play(1, 0, 62, 0.5, 0)), ("sched_cmd", 3, dict(clip=1, midi_channel=0, midi_note=62, change_volume=1, volume=0.9)),
            ("sched_queue_start", 2, 1), ("sched_run", 12, nframes)]
    ops += [("sched_start", 120), ("sched_cmd", 10, play(3, 2, 64, 0.6, 0)), ("sched_queue_start", 0, 3), ("sched_run", 40, nframes)]
    ops += [("sched_queue_start", 2, 4), ("sched_cmd", 96, play(1, 5, 67, 0.4, 0)), ("sched_cmd", 96, dict(clip=1, midi_channel=5, midi_note=67, change_looping=1, looping=1)),
            ("sched_run", 90, nframes)]
    ops += [("sched_queue_stop", 2, 1), ("sched_bpm", 90), ("sched_enable", 4, 3, 0), ("sched_run", 30, nframes)]
    ops += [("sched_enable", 2, 3, 1), ("sched_queue_start", 3, 6), ("sched_cmd", 400, play(0, 7, 55, 0.9, 0)), ("sched_run", 60, nframes)]
    ops += [("sched_stop",), ("sched_run", 10, nframes)]
    return ops


def random_scheduler_ops(seed, nframes):
    """Forty random scheduler calls (commands at assorted delays, queued starts/stops, start/stop/tempo changes in any order - so
    stop() while paused, stop() with a full ring, start() twice - channel switches), each followed by a run of periods."""
    rng = np.random.default_rng(seed)
    ops = []
    for _ in range(40):
        k = rng.integers(0, 8)
        cmd = dict(clip=int(rng.integers(0, 4)), midi_channel=int(rng.integers(-2, 10)), midi_note=int(rng.choice([60, 62, 64])),
                   change_volume=int(rng.integers(0, 2)), volume=float(np.float32(rng.uniform(0, 1))), looping=int(rng.integers(0, 2)),
                   change_looping=int(rng.integers(0, 2)), start_playback=int(rng.integers(0, 2)), stop_playback=int(rng.integers(0, 2)))
        if rng.random() < 0.3:
            cmd.update(change_slice=1, slice=int(rng.integers(0, 4)))
        if k <= 2:
            ops.append(("sched_cmd", int(rng.choice([0, 1, 2, 7, 96, 383, 384, 1000])), cmd))
        elif k == 3:
            ops.append(("sched_queue_start", int(rng.integers(0, 4)), int(rng.integers(-1, 10))))
        elif k == 4:
            ops.append(("sched_queue_stop", int(rng.integers(0, 4)), int(rng.integers(-1, 10))))
        elif k == 5:
            ops.append((("sched_start", int(rng.choice([60, 120, 137, 200]))) if rng.random() < 0.7 else ("sched_stop",)))
        elif k == 6:
            ops.append(("sched_bpm", int(rng.choice([50, 90, 120, 174]))))
        else:
            ops.append(("sched_enable", int(rng.integers(0, 20)), int(rng.integers(-2, 10)), int(rng.integers(0, 2))))
        ops.append(("sched_run", int(rng.integers(1, 40)), nframes))
    return ops


def scenario_scheduler_random(seed, nframes=256):
    return [op for op in scenario_scheduler(nframes) if op[0] == "clip"] + random_scheduler_ops(seed, nframes)


# ---------------------------------------------------------------------------------------------------------------
SCENARIOS = {
    "single_one_shot": lambda: scenario_single_voice(),
    "single_position_loop": lambda: scenario_single_voice(looping=True, length_beats=1.5),
    "polyphonic_seed2": lambda: scenario_polyphonic(seed=2),
    "polyphonic_seed7_ragged": lambda: scenario_polyphonic(seed=7, nframes=300, nblocks=9),
    "sketch_beat_aligned": lambda: scenario_sketch(True),
    "sketch_position_loops": lambda: scenario_sketch(False, nblocks=60),
    "scheduler_sketch": lambda: scenario_scheduler(),
    "scheduler_random_seed3": lambda: scenario_scheduler_random(3),
    "scheduler_random_seed4_1024": lambda: scenario_scheduler_random(4, 1024),
    "scheduler_random_seed8_64": lambda: scenario_scheduler_random(8, 64),
}


def digest(results, skip_buses=None):
    """sha256 over everything compare() looks at bit for bit (audio planes of the enabled buses, integer meter peaks,
    positions), so a library can be checked against the reference without the reference: tests/golden/libzl_digests.json."""
    import hashlib
    h = hashlib.sha256()
    for i, r in enumerate(results):
        keep = np.ones(BUSES, bool)
        if skip_buses:
            keep[skip_buses[i]] = False
        for k in ("raw", "dry", "fx1", "fx2"):
            a = np.ascontiguousarray(r[k][keep])
            a = a + np.float32(0.0)                                  # -0.0 and +0.0 compare equal: hash them alike
            h.update(a.tobytes())
        for _, live in r["positions"]:
            h.update(np.asarray(live, np.float32).tobytes())
        for m in r.get("meters", []):
            h.update(np.asarray(m[:2], np.int64).tobytes())
        if "clocks" in r:
            h.update(np.asarray(r["clocks"], np.uint64).tobytes())
            for per in r["commands"]:
                h.update(np.asarray([len(per)], np.int64).tobytes())
                for c in per:
                    h.update(np.asarray(c[:10], np.int64).tobytes() + np.asarray(c[10:], np.float32).tobytes())
            h.update(np.asarray(r["state"][:3], np.int64).tobytes())
    return h.hexdigest()
