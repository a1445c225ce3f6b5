"""Loader for the CPU oracle libraries (TEST INFRASTRUCTURE).

Only tests/, bench.py (cpu_baseline / --impl reference) and __graft_entry__.smoke() import this.
The product package libzl_b200 never does.
"""
import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")          # -ffp-contract=off : exact checker
ORACLE_FMA_SO = os.path.join(HERE, "liboracle_fma.so")  # -O3 -mfma contraction on : timed CPU baseline
JUCEREF_SO = os.path.join(HERE, "_ref", "libzl_juceref.so")


def build(ref=True):
    subprocess.check_call(["make", "-s", "-C", HERE, "oracle"])
    if ref and os.path.isdir("/root/reference/tracktion_engine/modules/juce/modules"):
        subprocess.check_call(["make", "-s", "-j4", "-C", HERE, "ref"])


def load_oracle(fma=False):
    import sys
    sys.path.insert(0, os.path.dirname(HERE))
    from libzl_b200 import _capi
    path = ORACLE_FMA_SO if fma else ORACLE_SO
    if not os.path.exists(path):
        build(ref=False)
    lib = _capi.Lib(path, "zlo_")
    d = lib.dll
    d.zlo_set_threads.restype, d.zlo_set_threads.argtypes = C.c_int, [C.c_void_p, C.c_int]
    d.zlo_adsr_run.restype = C.c_int
    d.zlo_adsr_run.argtypes = [C.c_double, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, C.c_int,
                               C.POINTER(C.c_float)]
    d.zlo_round_to_int.restype, d.zlo_round_to_int.argtypes = C.c_int, [C.c_double]
    d.zlo_set_float.restype, d.zlo_set_float.argtypes = None, [C.c_float, C.c_float, C.POINTER(C.c_int8)]
    return lib


def load_juceref():
    """The reference's own compiled JUCE objects, or None when oracle/_ref was not built."""
    if not os.path.exists(JUCEREF_SO):
        return None
    d = C.CDLL(JUCEREF_SO)
    f32p = C.POINTER(C.c_float)
    d.ref_adsr_run.restype = C.c_int
    d.ref_adsr_run.argtypes = [C.c_double, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, C.c_int, f32p]
    d.ref_jvoice_render.restype = C.c_int
    d.ref_jvoice_render.argtypes = [C.POINTER(f32p), C.c_int, C.c_longlong, C.c_double, C.c_double, C.c_int, C.c_int,
                                    C.c_float, C.c_double, C.c_double, C.c_int, C.c_int, C.c_int, f32p, f32p]
    d.ref_read_max_levels.restype = C.c_int
    d.ref_read_max_levels.argtypes = [C.POINTER(f32p), C.c_int, C.c_longlong, C.c_double, C.c_longlong, C.c_longlong, f32p]
    d.ref_round_to_int.restype, d.ref_round_to_int.argtypes = C.c_int, [C.c_double]
    d.ref_round_to_int_f.restype, d.ref_round_to_int_f.argtypes = C.c_int, [C.c_float]
    d.ref_rms.restype, d.ref_rms.argtypes = C.c_float, [f32p, C.c_int]
    d.ref_magnitude.restype, d.ref_magnitude.argtypes = C.c_float, [f32p, C.c_int]
    d.ref_find_min_max.restype, d.ref_find_min_max.argtypes = None, [f32p, C.c_int, f32p, f32p]
    d.ref_wav_roundtrip.restype = C.c_longlong
    d.ref_wav_roundtrip.argtypes = [C.POINTER(f32p), C.c_int, C.c_longlong, C.c_double, C.c_int, C.POINTER(C.c_ubyte),
                                    C.c_longlong, C.POINTER(C.c_longlong), C.POINTER(f32p)]
    if hasattr(d, "ref_thumb_create"):          # the reference's own juce::AudioThumbnail (oracle/ref/juce_thumb_ref.cpp)
        u8p, i8p = C.POINTER(C.c_ubyte), C.POINTER(C.c_byte)
        d.ref_thumb_create.restype, d.ref_thumb_create.argtypes = C.c_void_p, [u8p, C.c_longlong, C.c_int]
        d.ref_thumb_destroy.restype, d.ref_thumb_destroy.argtypes = None, [C.c_void_p]
        d.ref_thumb_info.restype = C.c_int
        d.ref_thumb_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_double)]
        d.ref_thumb_level0.restype, d.ref_thumb_level0.argtypes = C.c_int, [C.c_void_p, i8p]
        d.ref_thumb_pixels.restype, d.ref_thumb_pixels.argtypes = C.c_int, [C.c_void_p, C.c_double, C.c_double, C.c_int, i8p]
        d.ref_thumb_draw.restype = C.c_int
        d.ref_thumb_draw.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_int, C.c_int, C.c_float, u8p]
        d.ref_thumb_save.restype, d.ref_thumb_save.argtypes = C.c_longlong, [C.c_void_p, u8p, C.c_longlong]
        d.ref_thumb_load.restype = C.c_int
        d.ref_thumb_load.argtypes = [u8p, C.c_longlong, C.c_int, i8p, C.c_longlong, C.POINTER(C.c_int)]
    if hasattr(d, "ref_wav_read"):
        d.ref_wav_read.restype = C.c_longlong
        d.ref_wav_read.argtypes = [C.POINTER(C.c_ubyte), C.c_longlong, C.c_longlong, C.POINTER(f32p), C.POINTER(C.c_double),
                                   C.POINTER(C.c_longlong)]
    return d
