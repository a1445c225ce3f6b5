"""The C++ boundary: include/zl_b200.h compiles as C and links; the reference-named facade classes (include/zl/*.h,
libzl_b200/host) compile, link and - on a GPU - drive BASELINE configs[0] plus a configs[2]-style track to the same bytes the
oracle gives through the C ABI; the host library exports the libzl.h symbols a ctypes host of the reference binds."""
import os
import shutil
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "libzl_b200")


def _build_cxx(tmp_path, name="facade_cfg1"):
    exe = str(tmp_path / name)
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "cxx", name + ".cpp"), "-o", exe, "-L" + PKG, "-l:libzl_b200_host.so",
                           "-l:libzl_b200.so", "-Wl,-rpath," + PKG])
    return exe


def test_c_header_compiles_as_c_and_links(tmp_path):
    src = tmp_path / "abi.c"
    src.write_text('#include "zl_b200.h"\n#include <stdio.h>\nint main(void) { zl_config c; zl_clip_command k; zl_wav_format f;\n'
                   '  if (zl_config_default(&c) != ZL_OK || zl_clip_command_default(&k) != ZL_OK) return 1;\n'
                   '  if (zl_wav_probe("RIFX", 4, &f) != ZL_ERR_INVALID) return 2;\n'
                   '  printf("%d %d %d\\n", c.buses_per_session, c.voices_per_bus, k.midi_note); return 0; }\n')
    exe = str(tmp_path / "abi")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-I" + os.path.join(ROOT, "include"), str(src), "-o", exe,
                           "-L" + PKG, "-l:libzl_b200.so", "-Wl,-rpath," + PKG])
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0 and out.stdout.split() == ["12", "8", "-1"]


def test_facade_compiles_links_and_fails_loudly_without_a_device(tmp_path):
    exe = _build_cxx(tmp_path)
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a device is present: covered by the GPU test")
    except ImportError:
        pass
    r = subprocess.run([exe, "a.wav", "b.wav", str(tmp_path / "o.bin")], capture_output=True, text=True)
    assert r.returncode == 3 and "no CUDA device" in r.stderr          # no CPU fallback
    r = subprocess.run([_build_cxx(tmp_path, "facade_sched"), "a.wav", "b.wav", str(tmp_path / "o.bin")], capture_output=True, text=True)
    assert r.returncode == 3 and "no CUDA device" in r.stderr


def test_host_library_exports_the_reference_c_symbols():
    nm = shutil.which("nm")
    syms = subprocess.run([nm, "-D", "--defined-only", os.path.join(PKG, "libzl_b200_host.so")], capture_output=True, text=True, check=True).stdout
    for name in ("ClipAudioSource_new", "ClipAudioSource_play", "ClipAudioSource_stopOnChannel", "ClipAudioSource_setADSRRelease",
                 "ClipAudioSource_getDuration", "JackPassthrough_setPanAmount", "JackPassthrough_getDryAmount", "initJuce", "stopClips",
                 "dBFromVolume"):                                     # lib/libzl.h:18-179, bound by test/playtest.py:25-47
        assert f" T {name}\n" in syms, name
    for cls in ("SamplerSynth", "ClipAudioSource", "JackPassthrough", "AudioLevels", "WaveFormItem", "SamplerSynthVoice", "SyncTimer"):
        assert cls in syms


@pytest.mark.gpu
def test_facade_cfg1_and_a_sketch_track_equal_the_oracle(tmp_path, oracle, product):
    from libzl_b200 import _capi as K
    from libzl_b200.engine import Engine
    import thumbref as T
    rng = np.random.default_rng(77)
    img_clip = T.wav_image(rng.uniform(-1, 1, (2, 9000)).astype(np.float32), 48000, 16)
    img_loop = T.wav_image(rng.uniform(-1, 1, (2, 48000)).astype(np.float32), 48000, -32)
    (tmp_path / "clip.wav").write_bytes(img_clip.tobytes())
    (tmp_path / "loop.wav").write_bytes(img_loop.tobytes())
    exe = _build_cxx(tmp_path)
    out = tmp_path / "out.bin"
    r = subprocess.run([exe, str(tmp_path / "clip.wav"), str(tmp_path / "loop.wav"), str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    got = np.frombuffer(out.read_bytes(), np.uint8)

    # the same session through the C ABI on the oracle
    e = Engine(lib=oracle, n_sessions=1, buses_per_session=12, voices_per_bus=8, sample_rate=48000.0, max_frames=512, max_blocks=8)
    c0, c1 = e.add_clip_wav(img_clip), e.add_clip_wav(img_loop)
    sl = [k / 16.0 for k in range(16)]
    e.set_clip_params(c0, slice_pos=sl)
    e.set_clip_params(c1, slice_pos=sl, attack=float(np.float32(0.002)), decay=float(np.float32(0.05)), sustain=float(np.float32(0.7)),
                      release=float(np.float32(0.03)), length_s=0.75, length_beats=1.5, pan=float(np.float32(-0.4)), volume_abs=float(np.float32(0.8)))
    e.set_passthrough(0, 5, dry=0.8, fx1=0.3, fx2=0.0, pan=0.25)
    e.command(0, 0, 0, clip=c0, midi_channel=-2, midi_note=60, change_volume=1, volume=1.0, looping=0, start_playback=1)
    e.command(0, 0, 0, clip=c1, midi_channel=3, midi_note=60, change_volume=1, volume=1.0, looping=1, stop_playback=1, start_playback=1)
    exp = []
    for rr in range(3):
        ck = (K.zl_block_clock * 8)()
        for b in range(8):
            cur = 1000000 + (rr * 8 + b) * 10667
            ck[b].current_usecs, ck[b].next_usecs, ck[b].playhead, ck[b].playhead_usecs, ck[b].subbeat_usecs = cur, cur + 10667, 0, 1000000, 5208
        if rr == 2:
            e.command(0, 0, 0, clip=c1, midi_channel=3, midi_note=60, stop_playback=1)
        e.render(8, 512, ck)
        for bus, which in ((0, K.ZL_BUS_RAW), (5, K.ZL_BUS_RAW), (5, K.ZL_BUS_DRY), (5, K.ZL_BUS_FX1)):
            l, r_ = e.bus(0, bus, which)
            exp += [l.tobytes(), r_.tobytes()]
        pb = e.levels_tick(0, 0, 7, is_playback=True)
        tr = None
        for bus in range(2, 12):
            x = e.levels_tick(0, bus, 7)
            if bus == 5:
                tr = x
        hold = np.float32(-200.0) if pb.hold_a <= 0 else max(np.float32(-200.0), np.float32(20) * np.log10(np.float32(pb.hold_a), dtype=np.float32))
        exp.append(np.array([tr.db_a, tr.db_b, tr.db_sum, pb.db_a, hold], np.float32).tobytes())
        g, p = e.positions(c1)
        exp.append(np.array([e.clip_peak_gain(c1), p[0] if len(p) else -1.0], np.float32).tobytes())
    dur = float(np.float32(np.float64(48000) / 48000.0))
    exp.append(e.thumb_raster(c1, 0, 0.0, dur, 200, 64, 1.0).tobytes())
    e.close()
    exp = np.frombuffer(b"".join(exp), np.uint8)
    assert got.size == exp.size
    # audio, masks and meters: byte for byte, except the five dB floats per render (libm log10f on two machines: 1e-4 dB)
    n_audio = 8 * 8 * 512 * 4
    per = n_audio + 5 * 4 + 2 * 4
    for rr in range(3):
        a, b = got[rr * per:(rr + 1) * per], exp[rr * per:(rr + 1) * per]
        assert np.array_equal(a[:n_audio], b[:n_audio]), f"render {rr}: ports differ"
        assert np.allclose(a[n_audio:n_audio + 20].view(np.float32), b[n_audio:n_audio + 20].view(np.float32), rtol=1e-6, atol=1e-4)
        assert np.array_equal(a[n_audio + 20:], b[n_audio + 20:]), f"render {rr}: positions model differs"
    assert np.array_equal(got[3 * per:], exp[3 * per:]) and got[3 * per:].sum() > 100
    assert np.abs(got[:n_audio].view(np.float32)).max() > 0.05


@pytest.mark.gpu
def test_facade_synctimer_session_equals_the_oracle(tmp_path, oracle, product):
    """tests/cxx/facade_sched.cpp (SyncTimer + SamplerSynth + ClipAudioSource, period by period and batched) against the
    same calls on the oracle's scheduler and renderer through the C ABI, one period per call: ports, clocks, command count."""
    import ctypes as C
    from libzl_b200 import _capi as K
    from libzl_b200.engine import Engine
    import thumbref as T
    rng = np.random.default_rng(78)
    img_clip = T.wav_image(rng.uniform(-1, 1, (2, 9000)).astype(np.float32), 48000, 16)
    img_loop = T.wav_image(rng.uniform(-1, 1, (2, 48000)).astype(np.float32), 48000, -32)
    (tmp_path / "clip.wav").write_bytes(img_clip.tobytes())
    (tmp_path / "loop.wav").write_bytes(img_loop.tobytes())
    out = tmp_path / "sched.bin"
    r = subprocess.run([_build_cxx(tmp_path, "facade_sched"), str(tmp_path / "clip.wav"), str(tmp_path / "loop.wav"), str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    got = np.frombuffer(out.read_bytes(), np.uint8)

    NF, PERIOD = 256, 5333
    e = Engine(lib=oracle, n_sessions=1, buses_per_session=12, voices_per_bus=8, sample_rate=48000.0, max_frames=NF, max_blocks=16)
    clip, loop = e.add_clip_wav(img_clip), e.add_clip_wav(img_loop)
    sl = [k / 16.0 for k in range(16)]
    e.set_clip_params(clip, slice_pos=sl)
    e.set_clip_params(loop, slice_pos=sl, attack=float(np.float32(0.002)), decay=float(np.float32(0.05)), sustain=float(np.float32(0.7)),
                      release=float(np.float32(0.03)), length_s=0.75, length_beats=1.5, volume_abs=float(np.float32(0.8)))
    s = C.c_void_p()
    assert oracle.sched_create(48000.0, 1024, C.byref(s)) == K.ZL_OK
    exp, state = [], {"period": 0, "sent": 0}

    def cmd(**f):
        c = K.zl_clip_command()
        oracle.clip_command_default(C.byref(c))
        for k, v in f.items():
            setattr(c, k, v)
        return c

    def play(cl, ch, note, vol, lp):
        return cmd(clip=cl, midi_channel=ch, midi_note=note, change_volume=1, volume=vol, looping=lp, stop_playback=lp, start_playback=1)

    def run(periods):
        for _ in range(periods):
            ck = (K.zl_block_clock * 1)()
            ck[0].current_usecs = 1_000_000 + state["period"] * PERIOD
            ck[0].next_usecs = ck[0].current_usecs + PERIOD
            ev = (K.zl_sched_event * 256)()
            n = C.c_int32()
            assert oracle.sched_advance(s, 1, NF, ck, ev, 256, C.byref(n)) == K.ZL_OK
            for i in range(n.value):
                if ev[i].kind == K.ZL_SCHED_CLIP_COMMAND:
                    e._chk(oracle.command(e.ctx, 0, 0, ev[i].tick, C.byref(ev[i].cmd)), "command")
                    state["sent"] += 1
                else:
                    e.set_bus_enabled(0, ev[i].channel + 2, bool(ev[i].enabled))
            e.render(1, NF, ck)
            state["period"] += 1
            yield

    def dump(nb, planes):
        for ch in (-2, 0, 1, 2, 3):
            exp.append(np.concatenate([p[ch + 2][0] for p in planes]).tobytes())
            exp.append(np.concatenate([p[ch + 2][1] for p in planes]).tobytes())
        st = [C.c_uint64() for _ in range(4)]
        oracle.sched_state(s, C.byref(st[0]), C.byref(st[1]), C.byref(st[2]), C.byref(st[3]), None, None)
        exp.append(np.array([x.value for x in st], np.uint64).tobytes())

    def periods(total, batch):
        done = 0
        while done < total:
            nb = min(batch, total - done)
            planes = []
            for _ in run(nb):
                planes.append([e.bus(0, b, K.ZL_BUS_RAW) for b in range(12)])
            dump(nb, planes)
            done += nb

    oracle.sched_schedule(s, C.byref(play(clip, -2, 60, 0.7, 0)), 0)
    oracle.sched_schedule(s, C.byref(play(loop, 0, 62, 0.5, 1)), 3)
    oracle.sched_schedule(s, C.byref(cmd(clip=loop, midi_channel=0, midi_note=62, change_volume=1, volume=0.9)), 3)
    oracle.sched_queue_start(s, loop, 1)
    periods(12, 1)
    oracle.sched_start(s, 120)
    oracle.sched_schedule(s, C.byref(play(clip, 2, 64, 0.6, 0)), 10)
    oracle.sched_queue_start(s, clip, 3)
    periods(40, 8)
    oracle.sched_set_bpm(s, 90)
    oracle.sched_schedule_channel_enabled(s, 4, 3, 0)
    oracle.sched_queue_stop(s, loop, 1)
    oracle.sched_schedule(s, C.byref(play(loop, 2, 67, 0.4, 0)), 96)
    periods(30, 1)
    oracle.sched_schedule(s, C.byref(play(clip, 0, 55, 0.9, 0)), 400)
    oracle.sched_stop(s)
    periods(10, 10)
    exp.append(np.array([state["sent"], state["sent"], state["period"]], np.uint64).tobytes())
    oracle.sched_destroy(s)
    e.close()
    exp = np.frombuffer(b"".join(exp), np.uint8)
    assert got.size == exp.size
    assert np.array_equal(got, exp), f"first difference at byte {int(np.argmax(got != exp))} of {got.size}"
    assert state["sent"] >= 8 and np.abs(got[:12 * 10 * NF * 4].view(np.float32)).max() > 0.05
