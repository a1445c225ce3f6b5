"""Host-side view of one zl context (include/zl_b200.h) with numpy buffers.

``Engine(lib=...)`` drives whichever implementation of the header it is given; the default is the
CUDA product.  tests/ pass the CPU oracle's library here to run *the same calls* on both sides.
"""
import ctypes as C
import os

import numpy as np

from . import _capi as K


class ZlError(RuntimeError):
    pass


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


# numpy view of zl_clip_command (include/zl_b200.h; lib/ClipCommand.h:11-31)
CMD_DTYPE = np.dtype([(n, np.float32 if t is C.c_float else np.int32) for n, t in K.zl_clip_command._fields_])
assert CMD_DTYPE.itemsize == C.sizeof(K.zl_clip_command)


def default_commands(n):
    """n commands with ClipCommand's defaults (clip -1, note -1, channel -1, slice -1)."""
    c = np.zeros(n, CMD_DTYPE)
    c["clip"] = -1; c["midi_note"] = -1; c["midi_channel"] = -1; c["slice"] = -1
    return c


class Engine:
    def __init__(self, lib=None, **cfg):
        if lib is None:
            from . import load
            lib = load()
        self.lib = lib
        c = K.zl_config()
        self._chk(lib.config_default(C.byref(c)), None)
        for k, v in cfg.items():
            if not hasattr(c, k):
                raise TypeError(f"unknown zl_config field {k!r}")
            setattr(c, k, v)
        self.cfg = c
        self.ctx = C.c_void_p()
        rc = lib.ctx_create(C.byref(c), C.byref(self.ctx))
        if rc != K.ZL_OK:
            raise ZlError(f"ctx_create failed with status {rc}"
                          + (" (no CUDA device: the product has no CPU fallback)" if rc == K.ZL_ERR_NODEVICE else ""))
        self.n = 0
        self._keep = []

    # -- plumbing -------------------------------------------------------------------------
    def _chk(self, rc, what="call"):
        if rc != K.ZL_OK:
            msg = ""
            if getattr(self, "ctx", None):
                e = self.lib.last_error(self.ctx)
                msg = e.decode() if e else ""
            raise ZlError(f"zl {what} failed: status {rc} {msg}")

    def close(self):
        if getattr(self, "ctx", None):
            self.lib.ctx_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def n_buses(self):
        return self.cfg.n_sessions * self.cfg.buses_per_session

    @property
    def n_voices(self):
        return self.n_buses * self.cfg.voices_per_bus

    def sync(self):
        self._chk(self.lib.sync(self.ctx), "sync")

    def set_stream(self, stream_ptr):
        self._chk(self.lib.ctx_set_stream(self.ctx, C.c_void_p(stream_ptr)), "ctx_set_stream")

    # -- clips ----------------------------------------------------------------------------
    def add_clip(self, pcm, rate=48000.0):
        """pcm: float32 [nch, frames] planar (ClipAudioSource_new + registerClip)."""
        pcm = np.ascontiguousarray(np.atleast_2d(pcm), dtype=np.float32)
        nch, n = pcm.shape
        ptrs = (C.POINTER(C.c_float) * nch)(*[_fp(pcm[i]) for i in range(nch)])
        cid = C.c_int32(-1)
        self._chk(self.lib.clip_create_from_pcm(self.ctx, ptrs, nch, n, float(rate), C.byref(cid)), "clip_create_from_pcm")
        return cid.value

    def wav_probe(self, wav):
        """What the reference's WAV reader would find in the image (host-only chunk walk); raises on images it refuses."""
        buf = np.frombuffer(bytes(wav), np.uint8) if not isinstance(wav, np.ndarray) else np.ascontiguousarray(wav, np.uint8)
        f = K.zl_wav_format()
        self._chk(self.lib.wav_probe(buf.ctypes.data_as(C.c_void_p), buf.size, C.byref(f)), "wav_probe")
        return f

    def add_clip_wav(self, wav):
        """ClipAudioSource_new(path) with the file already in memory: bytes of a WAV file -> HBM clip, decoded on the device."""
        buf = np.frombuffer(bytes(wav), np.uint8) if not isinstance(wav, np.ndarray) else np.ascontiguousarray(wav, np.uint8)
        cid = C.c_int32(-1)
        self._chk(self.lib.clip_create_from_wav(self.ctx, buf.ctypes.data_as(C.c_void_p), buf.size, C.byref(cid)), "clip_create_from_wav")
        return cid.value

    def add_clip_wav_file(self, path):
        cid = C.c_int32(-1)
        self._chk(self.lib.clip_create_from_wav_file(self.ctx, os.fsencode(path), C.byref(cid)), "clip_create_from_wav_file")
        return cid.value

    def clip_pcm(self, clip, nch, frames):
        """The clip's sample data as SamplerSynthSound holds it: float32 [nch, frames]."""
        out = np.zeros((nch, frames), np.float32)
        ptrs = (C.POINTER(C.c_float) * nch)(*[_fp(out[i]) for i in range(nch)])
        self._chk(self.lib.clip_read_pcm(self.ctx, clip, ptrs, nch, frames), "clip_read_pcm")
        return out

    def wav_encode(self, left, right=None, rate=48000.0, bits=16):
        """Host floats -> the bytes of the WAV file the reference's writer would produce (16 or 24 bit)."""
        left = np.ascontiguousarray(left, np.float32)
        right = None if right is None else np.ascontiguousarray(right, np.float32)
        size = C.c_uint64(0)
        cap = 104 + left.size * (2 if right is not None else 1) * (bits // 8) + 1
        out = np.zeros(cap, np.uint8)
        self._chk(self.lib.wav_encode(self.ctx, _fp(left), _fp(right) if right is not None else None, left.size, float(rate), bits,
                                      out.ctypes.data_as(C.c_void_p), cap, C.byref(size)), "wav_encode")
        return out[:size.value].copy()

    def bounce_wav(self, session=0, bus=-1, which=0, bits=16):
        """The last render of one stereo port (bus < 0: the session's master) as a WAV file image."""
        size = C.c_uint64(0)
        self.lib.bounce_wav(self.ctx, session, bus, which, bits, None, 0, C.byref(size))
        if size.value == 0:
            self._chk(K.ZL_ERR_STATE, "bounce_wav")
        out = np.zeros(size.value, np.uint8)
        self._chk(self.lib.bounce_wav(self.ctx, session, bus, which, bits, out.ctypes.data_as(C.c_void_p), out.size, C.byref(size)), "bounce_wav")
        return out

    def masters_pcm(self, bits=16):
        """Every session's master of the last render as PCM (zl_read_master_all_pcm_async + zl_read_wait):
        uint8 [n_sessions][frames * 2 * bits/8], each row the data chunk of that session's bounced WAV file."""
        n = self.nblocks * self.nframes
        rows = self.cfg.n_sessions
        nbytes = rows * n * 2 * (bits // 8)
        hp = C.c_void_p()
        self._chk(self.lib.host_alloc(C.byref(hp), max(nbytes, 16)), "host_alloc")
        try:
            self._chk(self.lib.read_master_all_pcm_async(self.ctx, hp, n, bits), "read_master_all_pcm_async")
            self._chk(self.lib.read_wait(self.ctx), "read_wait")
            return np.ctypeslib.as_array(C.cast(hp, C.POINTER(C.c_uint8)), shape=(rows, nbytes // rows)).copy()
        finally:
            self.lib.host_free(hp)

    def thumb_save_jatm(self, clip):
        """The clip's thumbnail as a JUCE `jatm` stream (AudioThumbnail::saveTo)."""
        size = C.c_uint64(0)
        self.lib.thumb_save_jatm(self.ctx, clip, None, 0, C.byref(size))
        if size.value == 0:
            self._chk(K.ZL_ERR_STATE, "thumb_save_jatm")
        out = np.zeros(size.value, np.uint8)
        self._chk(self.lib.thumb_save_jatm(self.ctx, clip, out.ctypes.data_as(C.c_void_p), out.size, C.byref(size)), "thumb_save_jatm")
        return out

    def thumb_load_jatm(self, clip, jatm):
        buf = np.ascontiguousarray(jatm, np.uint8)
        self._chk(self.lib.thumb_load_jatm(self.ctx, clip, buf.ctypes.data_as(C.c_void_p), buf.size), "thumb_load_jatm")

    def add_clip_device(self, dptr, nch, frames, rate=48000.0):
        cid = C.c_int32(-1)
        self._chk(self.lib.clip_create_from_device(self.ctx, C.c_void_p(dptr), nch, frames, float(rate), C.byref(cid)),
                  "clip_create_from_device")
        return cid.value

    def clip_params(self, clip):
        p = K.zl_clip_params()
        self._chk(self.lib.clip_get_params(self.ctx, clip, C.byref(p)), "clip_get_params")
        return p

    def set_clip_params(self, clip, slice_pos=None, **kw):
        p = self.clip_params(clip)
        p.slice_pos = None
        for k, v in kw.items():
            if not hasattr(p, k):
                raise TypeError(f"unknown zl_clip_params field {k!r}")
            setattr(p, k, v)
        if slice_pos is not None:
            arr = (C.c_double * len(slice_pos))(*slice_pos)
            p.slice_pos = C.cast(arr, C.POINTER(C.c_double))
            p.n_slices = len(slice_pos)
        self._chk(self.lib.clip_set_params(self.ctx, clip, C.byref(p)), "clip_set_params")

    def slice_bounds(self, clip, slice_=-1):
        a, b = C.c_int32(), C.c_int32()
        self._chk(self.lib.clip_slice_bounds(self.ctx, clip, slice_, C.byref(a), C.byref(b)), "clip_slice_bounds")
        return a.value, b.value

    # -- commands -------------------------------------------------------------------------
    def command(self, session=0, block=0, tick=0, **fields):
        c = K.zl_clip_command()
        self.lib.clip_command_default(C.byref(c))
        for k, v in fields.items():
            if not hasattr(c, k):
                raise TypeError(f"unknown zl_clip_command field {k!r}")
            setattr(c, k, v)
        self._chk(self.lib.command(self.ctx, session, block, tick, C.byref(c)), "command")

    def command_batch(self, sessions, blocks, ticks, cmds):
        """sessions/blocks: int32 arrays, ticks: uint64 array, cmds: numpy structured array with the layout
        of zl_clip_command (see CMD_DTYPE) — one call for a whole period's command stream."""
        n = len(sessions)
        sessions = np.ascontiguousarray(sessions, np.int32)
        blocks = np.ascontiguousarray(blocks, np.int32)
        ticks = np.ascontiguousarray(ticks, np.uint64)
        cmds = np.ascontiguousarray(cmds, CMD_DTYPE)
        assert len(blocks) == n and len(ticks) == n and len(cmds) == n
        self._chk(self.lib.command_batch(self.ctx, n, sessions.ctypes.data_as(C.POINTER(C.c_int32)),
                                         blocks.ctypes.data_as(C.POINTER(C.c_int32)),
                                         ticks.ctypes.data_as(C.POINTER(C.c_uint64)),
                                         cmds.ctypes.data_as(C.POINTER(K.zl_clip_command))), "command_batch")

    def play(self, clip, loop=False, midi_channel=-2, note=60, volume=1.0, slice_=None, session=0, block=0, tick=0):
        """ClipAudioSource::play (lib/ClipAudioSource.cpp:415-429)."""
        f = dict(clip=clip, midi_channel=midi_channel, midi_note=note, change_volume=1, volume=volume,
                 looping=int(loop), stop_playback=int(loop), start_playback=1)
        if slice_ is not None:
            f.update(change_slice=1, slice=slice_)
        self.command(session, block, tick, **f)

    def stop(self, clip, midi_channel=-2, note=60, slice_=None, session=0, block=0, tick=0):
        """ClipAudioSource::stop (lib/ClipAudioSource.cpp:431-437)."""
        f = dict(clip=clip, midi_channel=midi_channel, midi_note=note, stop_playback=1)
        if slice_ is not None:
            f.update(change_slice=1, slice=slice_)
        self.command(session, block, tick, **f)

    def set_bus_enabled(self, session, bus, enabled):
        self._chk(self.lib.bus_set_enabled(self.ctx, session, bus, int(enabled)), "bus_set_enabled")

    def set_passthrough(self, session, bus, dry=1.0, fx1=1.0, fx2=1.0, pan=0.0, muted=False):
        p = K.zl_passthrough_params(dry, fx1, fx2, pan, int(muted))
        self._chk(self.lib.passthrough_set(self.ctx, session, bus, C.byref(p)), "passthrough_set")

    # -- render ---------------------------------------------------------------------------
    @staticmethod
    def make_clocks(nblocks, nframes, rate=48000.0, t0_usecs=1_000_000, first_block=0, playhead=0,
                    playhead_usecs=None, subbeat_usecs=5208):
        """Synthetic JACK cycle times: block i starts at t0 + i*period (SURVEY.md §8d cfg3 clock)."""
        period = int(round(nframes * 1e6 / rate))
        arr = (K.zl_block_clock * nblocks)()
        for i in range(nblocks):
            cur = t0_usecs + (first_block + i) * period
            arr[i].current_usecs = cur
            arr[i].next_usecs = cur + period
            arr[i].playhead = playhead
            arr[i].playhead_usecs = t0_usecs if playhead_usecs is None else playhead_usecs
            arr[i].subbeat_usecs = subbeat_usecs
        return arr

    def render(self, nblocks, nframes, clocks=None):
        if clocks is None:
            clocks = self.make_clocks(nblocks, nframes, self.cfg.sample_rate)
        self._chk(self.lib.render(self.ctx, nblocks, nframes, clocks), "render")
        self.n = nblocks * nframes
        self.nblocks, self.nframes = nblocks, nframes

    def bus(self, session, bus, which=K.ZL_BUS_RAW):
        l = np.empty(self.n, np.float32)
        r = np.empty(self.n, np.float32)
        self._chk(self.lib.read_bus(self.ctx, session, bus, which, _fp(l), _fp(r), self.n), "read_bus")
        return l, r

    def master(self, session=0):
        l = np.empty(self.n, np.float32)
        r = np.empty(self.n, np.float32)
        self._chk(self.lib.read_master(self.ctx, session, _fp(l), _fp(r), self.n), "read_master")
        return l, r

    def master_all(self):
        out = np.empty((self.cfg.n_sessions, 2, self.n), np.float32)
        self._chk(self.lib.read_master_all(self.ctx, _fp(out), self.n), "read_master_all")
        return out

    def block_levels(self):
        peaks = np.zeros((self.nblocks, self.n_buses, 2), np.int32)
        sumsq = np.zeros((self.nblocks, self.n_buses, 2), np.float64)
        self._chk(self.lib.read_block_levels(self.ctx, peaks.ctypes.data_as(C.POINTER(C.c_int32)),
                                             sumsq.ctypes.data_as(C.POINTER(C.c_double))), "read_block_levels")
        return peaks, sumsq

    def levels_tick(self, session, bus, block, is_playback=False):
        out = K.zl_levels()
        self._chk(self.lib.levels_tick(self.ctx, session, bus, block, int(is_playback), C.byref(out)), "levels_tick")
        return out

    def voices(self):
        arr = (K.zl_voice_info * self.n_voices)()
        self._chk(self.lib.read_voices(self.ctx, arr, self.n_voices), "read_voices")
        return arr

    def positions(self, clip):
        g = np.zeros(32, np.float32)
        p = np.zeros(32, np.float32)
        n = C.c_int32()
        self._chk(self.lib.read_positions(self.ctx, clip, _fp(g), _fp(p), 32, C.byref(n)), "read_positions")
        return g[:n.value], p[:n.value]

    def clip_peak_gain(self, clip):
        """ClipAudioSourcePositionsModel::peakGain (0.01 hysteresis, state kept per clip by the library)."""
        g = C.c_float()
        self._chk(self.lib.clip_peak_gain(self.ctx, clip, C.byref(g), None), "clip_peak_gain")
        return g.value

    def voice_frames(self):
        v = C.c_uint64()
        self._chk(self.lib.last_voice_frames(self.ctx, C.byref(v)), "last_voice_frames")
        return v.value

    def source_floats(self):
        v = C.c_uint64()
        self._chk(self.lib.last_source_floats(self.ctx, C.byref(v)), "last_source_floats")
        return v.value

    def total_counters(self):
        a, b = C.c_uint64(), C.c_uint64()
        self._chk(self.lib.total_counters(self.ctx, C.byref(a), C.byref(b)), "total_counters")
        return a.value, b.value

    def dropped_commands(self):
        v = C.c_uint64()
        self._chk(self.lib.dropped_commands(self.ctx, C.byref(v)), "dropped_commands")
        return v.value

    def launch_count(self):
        v = C.c_uint64()
        self._chk(self.lib.launch_count(self.ctx, C.byref(v)), "launch_count")
        return v.value

    def device_master(self):
        p, s = C.c_void_p(), C.c_int64()
        self._chk(self.lib.device_master_ptr(self.ctx, C.byref(p), C.byref(s)), "device_master_ptr")
        return p.value, s.value

    def device_bus(self, which):
        p, s = C.c_void_p(), C.c_int64()
        self._chk(self.lib.device_bus_ptr(self.ctx, which, C.byref(p), C.byref(s)), "device_bus_ptr")
        return p.value, s.value

    # -- thumbnails -----------------------------------------------------------------------
    def thumb_build(self, clip, samples_per_thumb=512):
        self._chk(self.lib.thumb_build(self.ctx, clip, samples_per_thumb), "thumb_build")

    def thumb_level0(self, clip):
        n, nch = C.c_int32(), C.c_int32()
        self._chk(self.lib.thumb_size(self.ctx, clip, C.byref(n), C.byref(nch)), "thumb_size")
        out = np.zeros((nch.value, n.value, 2), np.int8)
        self._chk(self.lib.thumb_read_level0(self.ctx, clip, out.ctypes.data_as(C.POINTER(C.c_int8))), "thumb_read_level0")
        return out

    def thumb_pixels(self, clip, t0, t1, width, nch=2):
        out = np.zeros((nch, max(width, 1), 2), np.int8)
        rc = self.lib.thumb_pixels(self.ctx, clip, float(t0), float(t1), width, out.ctypes.data_as(C.POINTER(C.c_int8)))
        if rc == K.ZL_ERR_INVALID:
            return None  # refillCache returned false
        self._chk(rc, "thumb_pixels")
        return out

    def thumb_raster(self, clip, channel, t0, t1, width, height, zoom=1.0):
        mask = np.zeros((height, width), np.uint8)
        rc = self.lib.thumb_raster(self.ctx, clip, channel, float(t0), float(t1), width, height, float(zoom),
                                   mask.ctypes.data_as(C.POINTER(C.c_uint8)))
        if rc == K.ZL_ERR_INVALID:
            return None
        self._chk(rc, "thumb_raster")
        return mask

    # -- standalone stages ------------------------------------------------------------------
    def passthrough_process(self, l, r, dry=1.0, fx1=1.0, fx2=1.0, pan=0.0, muted=False):
        l = np.ascontiguousarray(l, np.float32)
        r = np.ascontiguousarray(r, np.float32)
        out = np.zeros((6, l.size), np.float32)
        p = K.zl_passthrough_params(dry, fx1, fx2, pan, int(muted))
        self._chk(self.lib.passthrough_process(self.ctx, C.byref(p), _fp(l), _fp(r), l.size, _fp(out)), "passthrough_process")
        return out

    def levels_scan(self, l, r):
        l = np.ascontiguousarray(l, np.float32)
        r = np.ascontiguousarray(r, np.float32)
        pa, pb, sa, sb = C.c_int32(), C.c_int32(), C.c_double(), C.c_double()
        self._chk(self.lib.levels_scan(self.ctx, _fp(l), _fp(r), l.size, C.byref(pa), C.byref(pb), C.byref(sa), C.byref(sb)),
                  "levels_scan")
        return pa.value, pb.value, sa.value, sb.value
