"""Helpers for the thumbnail parity tests: WAV images and a Python view of the reference's own juce::AudioThumbnail
(oracle/ref/juce_thumb_ref.cpp inside oracle/_ref/libzl_juceref.so; TEST INFRASTRUCTURE)."""
import ctypes as C
import struct

import numpy as np

u8p, i8p = C.POINTER(C.c_ubyte), C.POINTER(C.c_byte)


def wav_image(pcm, rate=48000, bits=-32):
    """A canonical RIFF/WAVE image of planar float pcm [nch, n]: bits = -32 float32, 16 / 24 PCM (values rounded here,
    so the image - not the floats - is the test input)."""
    pcm = np.atleast_2d(pcm)
    nch, n = pcm.shape
    inter = np.ascontiguousarray(pcm.T)
    if bits == -32:
        data, fmt, bps = inter.astype("<f4").tobytes(), 3, 4
    elif bits == 16:
        data, fmt, bps = np.clip(np.rint(inter * 32767.0), -32768, 32767).astype("<i2").tobytes(), 1, 2
    elif bits == 24:
        v = np.clip(np.rint(inter.astype(np.float64) * 8388607.0), -8388608, 8388607).astype("<i4")
        data, fmt, bps = v.view(np.uint8).reshape(-1, 4)[:, :3].tobytes(), 1, 3
    else:
        raise ValueError(bits)
    hdr = (b"RIFF" + struct.pack("<I", 36 + len(data)) + b"WAVEfmt "
           + struct.pack("<IHHIIHH", 16, fmt, nch, int(rate), int(rate) * nch * bps, nch * bps, 8 * bps)
           + b"data" + struct.pack("<I", len(data)))
    return np.frombuffer(hdr + data, np.uint8).copy()


def valid_pixels(t0, t1, width, rate):
    """Zoomed-in branch of CachedWindow::refillCache (juce_AudioThumbnail.cpp:459-492): pixels whose first sample is
    negative are left untouched by the reference (stale cache contents); the boundary zero-fills them.  Returns the mask of
    the pixels the reference does write."""
    dt = (t1 - t0) / width
    t = t0
    s = int(np.rint(t0 * rate))
    ok = np.zeros(width, bool)
    for i in range(width):
        nxt = int(np.rint((t + dt) * rate))
        ok[i] = s >= 0
        t += dt
        s = nxt
    return ok


class RefThumb:
    """juce::AudioThumbnail over a WAV image, as WaveFormItem drives it."""

    def __init__(self, dll, image, spt=512):
        self.d = dll
        self.img = np.ascontiguousarray(image, np.uint8)
        self.h = dll.ref_thumb_create(self.img.ctypes.data_as(u8p), self.img.size, spt)
        if not self.h:
            raise RuntimeError("the reference's AudioThumbnail refused the image")
        nt, nch, tl = C.c_int(), C.c_int(), C.c_double()
        dll.ref_thumb_info(self.h, C.byref(nt), C.byref(nch), C.byref(tl))
        self.n_thumb, self.nch, self.total_length = nt.value, nch.value, tl.value

    def close(self):
        if self.h:
            self.d.ref_thumb_destroy(self.h)
            self.h = None

    def level0(self):
        out = np.zeros((self.nch, self.n_thumb, 2), np.int8)
        self.d.ref_thumb_level0(self.h, out.ctypes.data_as(i8p))
        return out

    def pixels(self, t0, t1, width):
        out = np.zeros((self.nch, max(width, 1), 2), np.int8)
        ok = self.d.ref_thumb_pixels(self.h, float(t0), float(t1), width, out.ctypes.data_as(i8p))
        return out if ok else None

    def raster(self, channel, t0, t1, width, height, zoom=1.0):
        mask = np.zeros((height, width), np.uint8)
        self.d.ref_thumb_draw(self.h, channel, float(t0), float(t1), width, height, float(zoom), mask.ctypes.data_as(u8p))
        return mask

    def save(self):
        n = self.d.ref_thumb_save(self.h, None, 0)
        out = np.zeros(n, np.uint8)
        self.d.ref_thumb_save(self.h, out.ctypes.data_as(u8p), n)
        return out


def ref_load_jatm(dll, jatm, spt=512):
    jatm = np.ascontiguousarray(jatm, np.uint8)
    nch = C.c_int()
    cap = jatm.size
    out = np.zeros(cap, np.int8)
    n = dll.ref_thumb_load(jatm.ctypes.data_as(u8p), jatm.size, spt, out.ctypes.data_as(i8p), cap, C.byref(nch))
    if n <= 0:
        return None
    return out[:nch.value * n * 2].reshape(nch.value, n, 2).copy()
