"""ctypes view of include/zl_b200.h.

The same declarations bind either library that implements the header: the product
(`libzl_b200/libzl_b200.so`, symbols ``zl_*``, CUDA sm_100a) or — from tests/ and bench.py only —
the CPU oracle (``zlo_*``).  Nothing in this package loads the oracle.
"""
import ctypes as C

ZL_OK = 0
ZL_ERR_INVALID, ZL_ERR_CUDA, ZL_ERR_NOMEM, ZL_ERR_STATE, ZL_ERR_NODEVICE = 1, 2, 3, 4, 5
ZL_MODE_Z, ZL_MODE_J = 0, 1
ZL_Q1_TAP_GAIN, ZL_Q2_PREINC, ZL_Q3_RENOTEOFF, ZL_QUIRKS_FAITHFUL = 1, 2, 4, 7
ZL_OUT_RAW, ZL_OUT_DRY, ZL_OUT_FX1, ZL_OUT_FX2 = 1, 2, 4, 8
ZL_OUT_MASTER, ZL_OUT_LEVELS, ZL_OUT_RMS, ZL_OUT_ALL = 16, 32, 64, 127
ZL_BUS_RAW, ZL_BUS_DRY, ZL_BUS_FX1, ZL_BUS_FX2 = 0, 1, 2, 3
ZL_SCHED_CLIP_COMMAND, ZL_SCHED_CHANNEL_ENABLED = 0, 1


class zl_config(C.Structure):
    _fields_ = [("sample_rate", C.c_double), ("n_sessions", C.c_int32), ("buses_per_session", C.c_int32),
                ("voices_per_bus", C.c_int32), ("max_clips", C.c_int32), ("max_slices", C.c_int32),
                ("max_frames", C.c_int32), ("max_blocks", C.c_int32), ("mode", C.c_int32),
                ("quirks", C.c_int32), ("outputs", C.c_int32), ("device", C.c_int32), ("reserved", C.c_int32)]


class zl_clip_params(C.Structure):
    _fields_ = [("start_s", C.c_float), ("length_s", C.c_float), ("length_beats", C.c_float), ("pan", C.c_float),
                ("volume_abs", C.c_float), ("attack", C.c_float), ("decay", C.c_float), ("sustain", C.c_float),
                ("release", C.c_float), ("root_note", C.c_int32), ("n_slices", C.c_int32), ("reserved", C.c_int32),
                ("slice_pos", C.POINTER(C.c_double))]


class zl_clip_command(C.Structure):
    _fields_ = [("clip", C.c_int32), ("midi_note", C.c_int32), ("midi_channel", C.c_int32),
                ("start_playback", C.c_int32), ("stop_playback", C.c_int32), ("change_slice", C.c_int32),
                ("slice", C.c_int32), ("change_looping", C.c_int32), ("looping", C.c_int32),
                ("change_pitch", C.c_int32), ("pitch_change", C.c_float), ("change_speed", C.c_int32),
                ("speed_ratio", C.c_float), ("change_gain_db", C.c_int32), ("gain_db", C.c_float),
                ("change_volume", C.c_int32), ("volume", C.c_float)]


class zl_block_clock(C.Structure):
    _fields_ = [("current_usecs", C.c_uint64), ("next_usecs", C.c_uint64), ("playhead", C.c_uint64),
                ("playhead_usecs", C.c_uint64), ("subbeat_usecs", C.c_uint64)]


class zl_sched_event(C.Structure):
    _fields_ = [("kind", C.c_int32), ("block", C.c_int32), ("tick", C.c_uint64), ("cmd", zl_clip_command),
                ("channel", C.c_int32), ("enabled", C.c_int32), ("reserved", C.c_int32)]


class zl_passthrough_params(C.Structure):
    _fields_ = [("dry_amount", C.c_float), ("wet_fx1_amount", C.c_float), ("wet_fx2_amount", C.c_float),
                ("pan_amount", C.c_float), ("muted", C.c_int32)]


class zl_voice_info(C.Structure):
    _fields_ = [("position", C.c_double), ("pitch_ratio", C.c_double), ("envelope", C.c_float), ("gain", C.c_float),
                ("pos_gain", C.c_float), ("pos_progress", C.c_float), ("playing", C.c_int32), ("clip", C.c_int32),
                ("adsr_state", C.c_int32), ("midi_note", C.c_int32), ("slice", C.c_int32), ("looping", C.c_int32)]


class zl_levels(C.Structure):
    _fields_ = [("peak_a", C.c_int32), ("peak_b", C.c_int32), ("db_a", C.c_float), ("db_b", C.c_float),
                ("db_sum", C.c_float), ("hold_a", C.c_float), ("hold_b", C.c_float), ("rms_a", C.c_float),
                ("rms_b", C.c_float), ("reserved", C.c_int32)]


class zl_wav_format(C.Structure):
    _fields_ = [("channels", C.c_int32), ("bits_per_sample", C.c_int32), ("is_float", C.c_int32), ("bytes_per_frame", C.c_int32),
                ("sample_rate", C.c_double), ("frames", C.c_int64), ("data_offset", C.c_uint64), ("data_bytes", C.c_uint64)]


_P = C.POINTER
_ctx = C.c_void_p
_f32p = _P(C.c_float)

# name -> (restype, argtypes): every symbol include/zl_b200.h declares.
SIGNATURES = {
    "ctx_create": (C.c_int, [_P(zl_config), _P(_ctx)]),
    "ctx_destroy": (C.c_int, [_ctx]),
    "last_error": (C.c_char_p, [_ctx]),
    "config_default": (C.c_int, [_P(zl_config)]),
    "ctx_set_stream": (C.c_int, [_ctx, C.c_void_p]),
    "ctx_set_overlap": (C.c_int, [_ctx, C.c_int32]),
    "ctx_set_graph": (C.c_int, [_ctx, C.c_int32]),
    "sync": (C.c_int, [_ctx]),
    "clip_create_from_pcm": (C.c_int, [_ctx, _P(_f32p), C.c_int32, C.c_int64, C.c_double, _P(C.c_int32)]),
    "clip_create_from_device": (C.c_int, [_ctx, C.c_void_p, C.c_int32, C.c_int64, C.c_double, _P(C.c_int32)]),
    "wav_probe": (C.c_int, [C.c_void_p, C.c_uint64, _P(zl_wav_format)]),
    "clip_create_from_wav": (C.c_int, [_ctx, C.c_void_p, C.c_uint64, _P(C.c_int32)]),
    "clip_create_from_wav_file": (C.c_int, [_ctx, C.c_char_p, _P(C.c_int32)]),
    "clip_read_pcm": (C.c_int, [_ctx, C.c_int32, _P(_f32p), C.c_int32, C.c_int64]),
    "wav_encode": (C.c_int, [_ctx, _f32p, _f32p, C.c_int64, C.c_double, C.c_int32, C.c_void_p, C.c_uint64, _P(C.c_uint64)]),
    "bounce_wav": (C.c_int, [_ctx, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_uint64, _P(C.c_uint64)]),
    "thumb_save_jatm": (C.c_int, [_ctx, C.c_int32, C.c_void_p, C.c_uint64, _P(C.c_uint64)]),
    "thumb_load_jatm": (C.c_int, [_ctx, C.c_int32, C.c_void_p, C.c_uint64]),
    "clip_destroy": (C.c_int, [_ctx, C.c_int32]),
    "clip_params_default": (C.c_int, [_ctx, C.c_int32, _P(zl_clip_params)]),
    "clip_set_params": (C.c_int, [_ctx, C.c_int32, _P(zl_clip_params)]),
    "clip_get_params": (C.c_int, [_ctx, C.c_int32, _P(zl_clip_params)]),
    "clip_slice_bounds": (C.c_int, [_ctx, C.c_int32, C.c_int32, _P(C.c_int32), _P(C.c_int32)]),
    "command": (C.c_int, [_ctx, C.c_int32, C.c_int32, C.c_uint64, _P(zl_clip_command)]),
    "clip_command_default": (C.c_int, [_P(zl_clip_command)]),
    "command_batch": (C.c_int, [_ctx, C.c_int32, _P(C.c_int32), _P(C.c_int32), _P(C.c_uint64), _P(zl_clip_command)]),
    "last_source_floats": (C.c_int, [_ctx, _P(C.c_uint64)]),
    "total_counters": (C.c_int, [_ctx, _P(C.c_uint64), _P(C.c_uint64)]),
    "bus_set_enabled": (C.c_int, [_ctx, C.c_int32, C.c_int32, C.c_int32]),
    "passthrough_set": (C.c_int, [_ctx, C.c_int32, C.c_int32, _P(zl_passthrough_params)]),
    "passthrough_get": (C.c_int, [_ctx, C.c_int32, C.c_int32, _P(zl_passthrough_params)]),
    "render": (C.c_int, [_ctx, C.c_int32, C.c_int32, _P(zl_block_clock)]),
    "read_bus": (C.c_int, [_ctx, C.c_int32, C.c_int32, C.c_int32, _f32p, _f32p, C.c_int64]),
    "read_master": (C.c_int, [_ctx, C.c_int32, _f32p, _f32p, C.c_int64]),
    "read_master_all": (C.c_int, [_ctx, _f32p, C.c_int64]),
    "read_master_all_async": (C.c_int, [_ctx, _f32p, C.c_int64]),
    "read_master_all_pcm_async": (C.c_int, [_ctx, C.c_void_p, C.c_int64, C.c_int32]),
    "read_wait": (C.c_int, [_ctx]),
    "device_bus_ptr": (C.c_int, [_ctx, C.c_int32, _P(C.c_void_p), _P(C.c_int64)]),
    "device_master_ptr": (C.c_int, [_ctx, _P(C.c_void_p), _P(C.c_int64)]),
    "read_block_levels": (C.c_int, [_ctx, _P(C.c_int32), _P(C.c_double)]),
    "levels_tick": (C.c_int, [_ctx, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _P(zl_levels)]),
    "read_voices": (C.c_int, [_ctx, _P(zl_voice_info), C.c_int32]),
    "read_positions": (C.c_int, [_ctx, C.c_int32, _f32p, _f32p, C.c_int32, _P(C.c_int32)]),
    "clip_peak_gain": (C.c_int, [_ctx, C.c_int32, _P(C.c_float), _P(C.c_double)]),
    "host_profile": (C.c_int, [_ctx, _P(C.c_double), _P(C.c_double), _P(C.c_uint64)]),
    "sched_create": (C.c_int, [C.c_double, C.c_int32, _P(C.c_void_p)]),
    "sched_destroy": (C.c_int, [C.c_void_p]),
    "sched_start": (C.c_int, [C.c_void_p, C.c_int32]),
    "sched_stop": (C.c_int, [C.c_void_p]),
    "sched_set_bpm": (C.c_int, [C.c_void_p, C.c_int32]),
    "sched_schedule": (C.c_int, [C.c_void_p, _P(zl_clip_command), C.c_uint64]),
    "sched_schedule_channel_enabled": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int32, C.c_int32]),
    "sched_queue_start": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32]),
    "sched_queue_stop": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32]),
    "sched_advance": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, _P(zl_block_clock), _P(zl_sched_event), C.c_int32, _P(C.c_int32)]),
    "sched_state": (C.c_int, [C.c_void_p, _P(C.c_uint64), _P(C.c_uint64), _P(C.c_uint64), _P(C.c_uint64), _P(C.c_int32), _P(C.c_uint64)]),
    "comm_unique_id": (C.c_int, [C.c_void_p]),
    "comm_init": (C.c_int, [_ctx, C.c_void_p, C.c_int32, C.c_int32]),
    "mix_reduce": (C.c_int, [_ctx, C.c_int32]),
    "comm_destroy": (C.c_int, [_ctx]),
    "last_voice_frames": (C.c_int, [_ctx, _P(C.c_uint64)]),
    "dropped_commands": (C.c_int, [_ctx, _P(C.c_uint64)]),
    "launch_count": (C.c_int, [_ctx, _P(C.c_uint64)]),
    "kernel_timing": (C.c_int, [_ctx, C.c_int32]),
    "kernel_time": (C.c_int, [_ctx, _P(C.c_double), _P(C.c_uint64)]),
    "thumb_build": (C.c_int, [_ctx, C.c_int32, C.c_int32]),
    "thumb_size": (C.c_int, [_ctx, C.c_int32, _P(C.c_int32), _P(C.c_int32)]),
    "thumb_read_level0": (C.c_int, [_ctx, C.c_int32, _P(C.c_int8)]),
    "thumb_pixels": (C.c_int, [_ctx, C.c_int32, C.c_double, C.c_double, C.c_int32, _P(C.c_int8)]),
    "thumb_raster": (C.c_int, [_ctx, C.c_int32, C.c_int32, C.c_double, C.c_double, C.c_int32, C.c_int32,
                               C.c_float, _P(C.c_uint8)]),
    "passthrough_process": (C.c_int, [_ctx, _P(zl_passthrough_params), _f32p, _f32p, C.c_int64, _f32p]),
    "levels_scan": (C.c_int, [_ctx, _f32p, _f32p, C.c_int64, _P(C.c_int32), _P(C.c_int32), _P(C.c_double),
                              _P(C.c_double)]),
    "levels_scan_static": (C.c_int, [_ctx, _f32p, _f32p, C.c_int64, _P(C.c_int32), _P(C.c_int32), _P(C.c_double),
                              _P(C.c_double)]),
    "host_alloc": (C.c_int, [_P(C.c_void_p), C.c_size_t]),
    "host_free": (C.c_int, [C.c_void_p]),
}


class Lib:
    """A loaded implementation of include/zl_b200.h with its symbol prefix stripped."""

    def __init__(self, path, prefix):
        self.path, self.prefix = path, prefix
        self.dll = C.CDLL(path)
        missing = []
        for name, (res, args) in SIGNATURES.items():
            try:
                fn = getattr(self.dll, prefix + name)
            except AttributeError:
                missing.append(prefix + name)
                continue
            fn.restype, fn.argtypes = res, args
            setattr(self, name, fn)
        if missing:
            raise ImportError(f"{path} does not export: {', '.join(missing)}")
