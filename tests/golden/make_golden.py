"""Generates tests/golden/juce_golden.npz from the REFERENCE'S OWN compiled JUCE objects
(oracle/_ref/libzl_juceref.so, built by `make -C oracle ref` from /root/reference in place).

Run in the build container only (needs /root/reference):  python tests/golden/make_golden.py
The vectors pin oracle/zl_oracle.cpp's restatements of juce::ADSR, juce::SamplerVoice::renderNextBlock,
AudioFormatReader::readMaxLevels, juce::roundToInt, AudioBuffer::getRMSLevel/getMagnitude and the WAV
int16/24 decode scale; the GPU box replays them without the reference.
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import loader  # noqa: E402

f32p = C.POINTER(C.c_float)


def fp(a):
    return a.ctypes.data_as(f32p)


def planar_ptrs(pcm):
    return (f32p * pcm.shape[0])(*[fp(pcm[i]) for i in range(pcm.shape[0])])


def main():
    ref = loader.load_juceref()
    if ref is None:
        loader.build(ref=True)
        ref = loader.load_juceref()
    out = {}
    rng = np.random.default_rng(20261104)

    # --- juce::ADSR  (juce_ADSR.h) : params (sr, a, d, s, r, n_on, n_total, renote)
    adsr_cases = [
        (48000.0, 0.0, 0.1, 1.0, 0.05, 600, 4000, 0),      # libzl ctor defaults: straight to sustain
        (44100.0, 0.1, 0.1, 0.5, 0.1, 12000, 20000, 0),    # ADSRTests parameters
        (48000.0, 0.01, 0.02, 0.7, 0.03, 2000, 5000, 0),
        (48000.0, 0.01, 0.02, 0.7, 0.03, 2000, 5000, 1),   # libzl Q3: noteOff re-armed on every release frame
        (48000.0, 0.005, 0.0, 0.6, 0.0, 700, 1200, 0),     # zero decay, zero release
        (22050.0, 0.0, 0.05, 0.25, 0.2, 100, 6000, 0),     # noteOff during decay
        (48000.0, 0.05, 0.1, 0.3, 0.01, 300, 2000, 1),     # noteOff during attack, renote
    ]
    out["adsr_params"] = np.array(adsr_cases, np.float64)
    for i, (sr, a, d, s, r, n_on, n_total, ren) in enumerate(adsr_cases):
        buf = np.zeros(int(n_total), np.float32)
        ref.ref_adsr_run(sr, a, d, s, r, int(n_on), int(n_total), int(ren), fp(buf))
        out[f"adsr_{i}"] = buf

    # --- juce::SamplerVoice::renderNextBlock through juce::Synthesiser
    # (nch, len, rate, outRate, root, note, velocity, attack, release, nblocks, nframes, noteoff_block)
    j_cases = [
        (2, 6000, 48000.0, 48000.0, 60, 60, 1.0, 0.0, 0.05, 8, 256, -1),
        (1, 5000, 48000.0, 48000.0, 60, 67, 0.8, 0.0, 0.05, 8, 256, -1),
        (2, 6000, 44100.0, 48000.0, 60, 53, 0.5, 0.01, 0.02, 10, 512, 3),
        (2, 3000, 48000.0, 48000.0, 64, 76, 0.9, 0.002, 0.01, 6, 512, 1),
        (1, 900, 48000.0, 44100.0, 60, 60, 1.0, 0.0, 0.1, 4, 300, -1),
    ]
    out["j_params"] = np.array(j_cases, np.float64)
    for i, (nch, n, rate, orate, root, note, vel, att, rel, nb, nf, off) in enumerate(j_cases):
        pcm = rng.uniform(-1, 1, (nch, n)).astype(np.float32)
        L = np.zeros(nb * nf, np.float32)
        R = np.zeros(nb * nf, np.float32)
        ref.ref_jvoice_render(planar_ptrs(pcm), nch, n, rate, orate, root, note, vel, att, rel, nb, nf, off, fp(L), fp(R))
        out[f"j_pcm_{i}"], out[f"j_L_{i}"], out[f"j_R_{i}"] = pcm, L, R

    # --- AudioFormatReader::readMaxLevels + findMinAndMax over 512-sample blocks
    pcm = (rng.uniform(-1, 1, (2, 512 * 40 + 137)) * np.linspace(0.0, 1.3, 512 * 40 + 137)).astype(np.float32)
    pcm[0, 1000] = 1.0; pcm[1, 2000] = -1.0; pcm[0, 512 * 7:512 * 8] = 0.25; pcm[1, 512 * 9:512 * 10] = 0.0
    lv = np.zeros((40, 4), np.float32)
    for b in range(40):
        ref.ref_read_max_levels(planar_ptrs(pcm), 2, pcm.shape[1], 48000.0, b * 512, 512, fp(lv[b]))
    out["rml_pcm"], out["rml_levels"] = pcm, lv
    odd = np.zeros((6, 4), np.float32)   # ragged ranges: (start, num)
    ranges = [(0, 1), (5, 3), (20000, 700), (20400, 512), (pcm.shape[1] - 10, 512), (777, 4097)]
    for k, (s, n) in enumerate(ranges):
        ref.ref_read_max_levels(planar_ptrs(pcm), 2, pcm.shape[1], 48000.0, s, n, fp(odd[k]))
    out["rml_ranges"], out["rml_ragged"] = np.array(ranges, np.int64), odd

    # --- juce::roundToInt (double and float overloads)
    xs = np.concatenate([np.arange(-130, 131) * 0.5, rng.uniform(-200, 200, 200), [126.5, 127.5, -127.5, -128.5, 1e-9, -1e-9]])
    out["rti_x"] = xs
    out["rti_d"] = np.array([ref.ref_round_to_int(float(x)) for x in xs], np.int32)
    xf = xs.astype(np.float32)
    out["rti_f"] = np.array([ref.ref_round_to_int_f(float(x)) for x in xf], np.int32)

    # --- getRMSLevel / getMagnitude
    sig = rng.uniform(-1, 1, 512).astype(np.float32)
    out["rms_x"] = sig
    out["rms"] = np.array([ref.ref_rms(fp(sig), 512), ref.ref_magnitude(fp(sig), 512)], np.float32)

    # --- WAV int16 / int24 decode through the reference's WavAudioFormat (for the ingest row n1)
    wpcm = rng.uniform(-1, 1, (2, 1000)).astype(np.float32)
    for bits in (16, 24):
        back = np.zeros((2, 1000), np.float32)
        bp = (f32p * 2)(fp(back[0]), fp(back[1]))
        cap = 1 << 16
        wav = np.zeros(cap, np.uint8)
        size = C.c_longlong()
        n = ref.ref_wav_roundtrip(planar_ptrs(wpcm), 2, 1000, 48000.0, bits, wav.ctypes.data_as(C.POINTER(C.c_ubyte)), cap, C.byref(size), bp)
        assert n == 1000, n
        out[f"wav{bits}_bytes"] = wav[:size.value].copy()
        out[f"wav{bits}_back"] = back
    out["wav_src"] = wpcm

    # --- row n1: WAV images through the reference's reader the way SamplerSynthSound::loadSoundData reads them
    # (ref_wav_read); the first four are written by the reference's own writer, the rest are hand-built to reach the
    # reader's other branches (int32 PCM, WAVE_FORMAT_EXTENSIBLE, chunks before data, a data chunk longer than the file,
    # odd-sized chunks, float specials).  Images the reader refuses are kept too (wavbad_*).
    import struct

    def juce_written(pcm, bits):
        nch, n = pcm.shape
        back = np.zeros((nch, n), np.float32)
        bp = (f32p * nch)(*[fp(back[i]) for i in range(nch)])
        cap = 1 << 18
        wav = np.zeros(cap, np.uint8)
        size = C.c_longlong()
        got = ref.ref_wav_roundtrip(planar_ptrs(pcm), nch, n, 44100.0, bits, wav.ctypes.data_as(C.POINTER(C.c_ubyte)), cap, C.byref(size), bp)
        assert got == n, got
        return wav[:size.value].copy()

    def riff(chunks):
        body = b"WAVE" + b"".join(t + struct.pack("<I", len(d) if ln is None else ln) + d + (b"\0" if len(d) & 1 else b"") for t, d, ln in chunks)
        return np.frombuffer(b"RIFF" + struct.pack("<I", len(body)) + body, np.uint8).copy()

    def fmt_chunk(tag, nch, rate, bits, ext_sub=None):
        bpf = nch * bits // 8
        d = struct.pack("<HHIIHH", tag, nch, rate, bpf * rate, bpf, bits)
        if ext_sub is not None:
            d += struct.pack("<HHI", 22, bits, 0) + ext_sub
        return d

    pcm_guid = bytes([1, 0, 0, 0, 0, 0, 0x10, 0, 0x80, 0, 0, 0xaa, 0, 0x38, 0x9b, 0x71])
    flt_guid = bytes([3]) + pcm_guid[1:]
    i32 = rng.integers(-2**31, 2**31, (700, 2), dtype=np.int64).astype(np.int32)
    i32[0] = (-2**31, 2**31 - 1); i32[1] = (1, -1); i32[2] = (0x7fffff7f, 0x01000001); i32[3] = (-0x7fffff81, 0x00ffffff)
    s16 = rng.integers(-32768, 32768, (1000, 2)).astype(np.int16)
    s16[0] = (-32768, 32767)
    f32 = rng.uniform(-1.5, 1.5, 300).astype(np.float32)
    f32[:8] = np.array([np.nan, np.inf, -np.inf, -0.0, 1e-42, -1e-45, 3.4e38, 1.0], np.float32)
    s24 = rng.integers(0, 256, 3 * 501).astype(np.uint8)
    s24[:6] = (0, 0, 0x80, 0xff, 0xff, 0x7f)
    images = [
        juce_written(rng.uniform(-1, 1, (2, 900)).astype(np.float32), 8),
        juce_written(rng.uniform(-1, 1, (1, 777)).astype(np.float32), 16),
        juce_written(rng.uniform(-1, 1, (3, 640)).astype(np.float32), 24),
        juce_written(rng.uniform(-2, 2, (2, 512)).astype(np.float32), -32),
        riff([(b"fmt ", fmt_chunk(1, 2, 48000, 32), None), (b"data", i32.tobytes(), None)]),
        riff([(b"fmt ", fmt_chunk(1, 2, 22050, 16), None), (b"LIST", b"INFOISFT\x05\0\0\0zl_b\0\0", None),
              (b"data", s16[:601].tobytes()[:-2], 4000)]),                      # promises 1000 frames, holds 600.5
        riff([(b"fmt ", fmt_chunk(0xfffe, 1, 96000, 32, flt_guid), None), (b"data", f32.tobytes(), None)]),
        riff([(b"fmt ", fmt_chunk(0xfffe, 1, 48000, 24, pcm_guid), None), (b"data", s24.tobytes(), None), (b"junk", b"abc", None)]),
        riff([(b"JUNK", bytes(28), None), (b"fmt ", fmt_chunk(3, 2, 48000, 32), None), (b"data", f32[:298].tobytes(), None)]),
    ]
    out["wavc_count"] = np.array([len(images)], np.int32)
    for i, img in enumerate(images):
        back = np.zeros((2, 4096), np.float32)
        bp = (f32p * 2)(fp(back[0]), fp(back[1]))
        rate = C.c_double()
        info = (C.c_longlong * 4)()
        n = ref.ref_wav_read(img.ctypes.data_as(C.POINTER(C.c_ubyte)), len(img), 4096, bp, C.byref(rate), info)
        assert n > 0, (i, n)
        keep = min(2, info[0])
        out[f"wavc_{i}_bytes"] = img
        out[f"wavc_{i}_back"] = back[:keep, :n].copy()
        out[f"wavc_{i}_info"] = np.array([info[0], info[1], info[2], info[3], rate.value], np.float64)
    bad = [
        riff([(b"fmt ", fmt_chunk(2, 2, 48000, 4), None), (b"data", bytes(64), None)]),       # ADPCM
        riff([(b"fmt ", fmt_chunk(3, 1, 48000, 64), None), (b"data", bytes(64), None)]),      # float64
        np.frombuffer(b"RIFX" + bytes(60), np.uint8).copy(),
        riff([(b"data", bytes(64), None)]),                                                    # no fmt chunk
    ]
    out["wavbad_count"] = np.array([len(bad)], np.int32)
    for i, img in enumerate(bad):
        back = np.zeros((2, 64), np.float32)
        bp = (f32p * 2)(fp(back[0]), fp(back[1]))
        n = ref.ref_wav_read(img.ctypes.data_as(C.POINTER(C.c_ubyte)), len(img), 64, bp, None, None)
        assert n == -2, (i, n)
        out[f"wavbad_{i}_bytes"] = img

    path = os.path.join(HERE, "juce_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
