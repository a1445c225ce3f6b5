"""Rows n1 / n3: WAV ingest and bounce.

The golden images and their decoded samples come from the reference's own compiled WavAudioFormat reader and writer
(tests/golden/make_golden.py, ref_wav_roundtrip / ref_wav_read), so this row's parity is pinned: the oracle restatement
is checked against them on the CPU, the CUDA path against both on the GPU — bit for bit (samples) and byte for byte (files).
"""
import ctypes as C
import struct

import numpy as np
import pytest

from libzl_b200 import _capi as K
from libzl_b200.engine import Engine, ZlError
import scenarios as S


def images(golden):
    return [(i, golden[f"wavc_{i}_bytes"], golden[f"wavc_{i}_back"], golden[f"wavc_{i}_info"]) for i in range(int(golden["wavc_count"][0]))]


def bad_images(golden):
    return [golden[f"wavbad_{i}_bytes"] for i in range(int(golden["wavbad_count"][0]))]


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def decode(lib, img):
    e = Engine(lib=lib)
    try:
        f = e.wav_probe(img)
        c = e.add_clip_wav(img)
        return f, e.clip_pcm(c, min(2, f.channels), f.frames)
    finally:
        e.close()


def pcm16_image(s16, rate=48000):
    """A plain 44-byte-header WAV around int16 frames [n, nch] (what most files on disk look like)."""
    n, nch = s16.shape
    data = np.ascontiguousarray(s16, "<i2").tobytes()
    hdr = b"RIFF" + struct.pack("<I", 36 + len(data)) + b"WAVEfmt " + struct.pack("<IHHIIHH", 16, 1, nch, rate, rate * nch * 2, nch * 2, 16)
    return np.frombuffer(hdr + b"data" + struct.pack("<I", len(data)) + data, np.uint8)


# ---- CPU: the oracle against the reference's own reader and writer ----------------------------------------------------
def test_oracle_decodes_like_the_reference_reader(oracle, golden):
    for i, img, back, info in images(golden):
        f, got = decode(oracle, img)
        assert (f.channels, f.bits_per_sample, f.is_float, f.frames, f.sample_rate) == tuple(info), i
        assert got.shape == back.shape, i
        assert np.array_equal(bits(got), bits(back)), f"image {i}: decoded samples differ from the reference reader's"


def test_oracle_decodes_the_writer_round_trip(oracle, golden):
    for b in (16, 24):
        _, got = decode(oracle, golden[f"wav{b}_bytes"])
        assert np.array_equal(bits(got), bits(golden[f"wav{b}_back"]))


def test_oracle_refuses_what_the_reference_refuses(oracle, golden):
    for img in bad_images(golden):
        e = Engine(lib=oracle)
        with pytest.raises(ZlError):
            e.add_clip_wav(img)
        e.close()


def test_oracle_writer_bytes_equal_the_reference_writer(oracle, golden):
    src = golden["wav_src"]
    e = Engine(lib=oracle)
    for b in (16, 24):
        got = e.wav_encode(src[0], src[1], rate=48000.0, bits=b)
        assert np.array_equal(got, golden[f"wav{b}_bytes"]), f"{b}-bit file differs from the reference writer's"
    e.close()


def test_product_probe_needs_no_device_and_matches_the_reference(golden):
    """zl_wav_probe is the host half of the ingest: callable on a box without a GPU, same answers as the reference reader."""
    import libzl_b200
    lib = libzl_b200.load()
    for i, img, back, info in images(golden):
        f = K.zl_wav_format()
        assert lib.wav_probe(img.ctypes.data_as(C.c_void_p), img.size, C.byref(f)) == K.ZL_OK, i
        assert (f.channels, f.bits_per_sample, f.is_float, f.frames, f.sample_rate) == tuple(info), i
    for img in bad_images(golden):
        f = K.zl_wav_format()
        assert lib.wav_probe(img.ctypes.data_as(C.c_void_p), img.size, C.byref(f)) == K.ZL_ERR_INVALID
    f = K.zl_wav_format()
    assert lib.wav_probe(None, 0, C.byref(f)) == K.ZL_ERR_INVALID


def test_probe_agrees_between_product_and_oracle_on_mangled_headers(oracle, golden):
    """Truncations and byte flips of a good header: both parsers must make the same call and report the same fields."""
    import libzl_b200
    lib = libzl_b200.load()
    rng = np.random.default_rng(7)
    base = [golden["wavc_5_bytes"], golden["wavc_7_bytes"], golden["wav16_bytes"]]
    for img in base:
        cases = [img[:k].copy() for k in (0, 3, 8, 12, 20, 36, 44, 60, 100, 130)]
        for _ in range(60):
            m = img.copy()
            m[rng.integers(0, min(len(m), 120), 2)] = rng.integers(0, 256, 2)
            cases.append(m)
        for m in cases:
            m = np.ascontiguousarray(m)
            fa, fb = K.zl_wav_format(), K.zl_wav_format()
            ra = lib.wav_probe(m.ctypes.data_as(C.c_void_p), m.size, C.byref(fa))
            rb = oracle.wav_probe(m.ctypes.data_as(C.c_void_p), m.size, C.byref(fb))
            assert ra == rb
            if ra == K.ZL_OK:
                assert bytes(fa) == bytes(fb)


# ---- GPU: the CUDA path ---------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_device_decode_matches_the_reference_reader(product, oracle, golden):
    for i, img, back, info in images(golden):
        f, got = decode(product, img)
        assert (f.channels, f.bits_per_sample, f.is_float, f.frames, f.sample_rate) == tuple(info), i
        assert np.array_equal(bits(got), bits(back)), f"image {i}: device decode differs from the reference reader's"
        _, exp = decode(oracle, img)
        assert np.array_equal(bits(got), bits(exp))
    for b in (16, 24):
        _, got = decode(product, golden[f"wav{b}_bytes"])
        assert np.array_equal(bits(got), bits(golden[f"wav{b}_back"]))


@pytest.mark.gpu
def test_device_refuses_bad_images(product, golden):
    for img in bad_images(golden):
        e = Engine(lib=product)
        with pytest.raises(ZlError):
            e.add_clip_wav(img)
        e.close()


@pytest.mark.gpu
@pytest.mark.parametrize("frames,nch", [(1, 2), (2047, 1), (2048, 2), (2049, 2), ((1 << 22) + 77, 2)])
def test_device_decode_int16_at_size(product, frames, nch):
    """Sizes around the CTA tile and a 4 M-frame file: f32(s16 << 16) * 2^-31 == s16 / 32768 exactly (numpy restatement)."""
    rng = np.random.default_rng(frames)
    s16 = rng.integers(-32768, 32768, (frames, nch)).astype(np.int16)
    s16[0] = -32768
    f, got = decode(product, pcm16_image(s16))
    assert (f.frames, f.channels) == (frames, nch)
    exp = (s16.T.astype(np.float32) / np.float32(32768.0))
    assert np.array_equal(bits(got), bits(exp))


@pytest.mark.gpu
def test_device_decode_24bit_and_float_at_size(product, oracle):
    rng = np.random.default_rng(24)
    n = 300001
    raw = rng.integers(0, 256, n * 2 * 3).astype(np.uint8)
    body = b"WAVEfmt " + struct.pack("<IHHIIHH", 16, 1, 2, 44100, 44100 * 6, 6, 24) + b"data" + struct.pack("<I", raw.size) + raw.tobytes()
    img = np.frombuffer(b"RIFF" + struct.pack("<I", len(body)) + body, np.uint8)
    _, got = decode(product, img)
    _, exp = decode(oracle, img)
    assert np.array_equal(bits(got), bits(exp))
    v = raw.reshape(n, 2, 3).astype(np.int64)
    i24 = (v[..., 0] | (v[..., 1] << 8) | (v[..., 2] << 16))
    i24 = np.where(i24 >= 1 << 23, i24 - (1 << 24), i24)
    assert np.array_equal(bits(got), bits((i24.T.astype(np.float32) / np.float32(8388608.0))))
    x = rng.standard_normal((n, 2)).astype(np.float32)
    body = b"WAVEfmt " + struct.pack("<IHHIIHH", 16, 3, 2, 48000, 48000 * 8, 8, 32) + b"data" + struct.pack("<I", x.nbytes) + x.tobytes()
    img = np.frombuffer(b"RIFF" + struct.pack("<I", len(body)) + body, np.uint8)
    _, got = decode(product, img)
    assert np.array_equal(bits(got), bits(x.T))


@pytest.mark.gpu
def test_device_writer_bytes_equal_the_reference_writer(product, oracle, golden):
    src = golden["wav_src"]
    e = Engine(lib=product)
    for b in (16, 24):
        got = e.wav_encode(src[0], src[1], rate=48000.0, bits=b)
        assert np.array_equal(got, golden[f"wav{b}_bytes"]), f"{b}-bit file differs from the reference writer's"
    # clamps, ties and a mono file at size, against the oracle restatement of the same writer
    rng = np.random.default_rng(3)
    x = rng.uniform(-1.2, 1.2, 1 << 20).astype(np.float32)
    x[:10] = [1.0, -1.0, 0.99999994, -0.99999994, 0.0, -0.0, 1.5, -7.0, 2.0 ** -17, -(2.0 ** -16) * 1.5]
    o = Engine(lib=oracle)
    for b in (16, 24):
        assert np.array_equal(e.wav_encode(x, x[::-1].copy(), rate=44100.0, bits=b), o.wav_encode(x, x[::-1].copy(), rate=44100.0, bits=b))
        assert np.array_equal(e.wav_encode(x[:99999], None, rate=22050.0, bits=b), o.wav_encode(x[:99999], None, rate=22050.0, bits=b))
    e.close(); o.close()


@pytest.mark.gpu
def test_wav_file_plays_like_the_pcm_clip_and_bounces_like_the_oracle(product, oracle, golden, tmp_path):
    """File on disk -> clip -> render -> bounce: every stage equal between the CUDA path and the oracle, and the played
    WAV clip equal to the same samples handed over as PCM."""
    img = golden["wavc_2_bytes"]                     # 24 bit, three channels: the first two are kept
    path = tmp_path / "clip.wav"
    path.write_bytes(img.tobytes())
    outs, files = [], []
    for lib in (product, oracle):
        e = Engine(lib=lib, mode=K.ZL_MODE_Z, quirks=K.ZL_QUIRKS_FAITHFUL)
        c = e.add_clip_wav_file(str(path))
        c2 = e.add_clip(golden["wavc_2_back"], rate=44100.0)
        e.play(c, note=67, volume=0.9)
        e.play(c2, note=67, volume=0.9, midi_channel=-1)
        e.render(4, 256, e.make_clocks(4, 256))
        outs.append(S.collect(e))
        files.append([e.bounce_wav(0, -1, 0, b) for b in (16, 24)] + [e.bounce_wav(0, 0, 1, 16)])
        e.close()
    assert not S.compare(outs[0], outs[1], exact_audio=True)
    raw = outs[0]["raw"]
    assert np.abs(raw).max() > 0.01
    assert np.array_equal(raw[0], raw[1]), "the WAV clip (bus 0) and the PCM clip (bus 1) must render identically"
    for a, b in zip(*files):
        assert np.array_equal(a, b)
    # and the bounced file is itself a WAV the ingest reads back: the master quantised to 16 bits
    f, back = decode(product, files[0][0])
    assert (f.channels, f.bits_per_sample, f.frames, f.sample_rate) == (2, 16, 4 * 256, 48000.0)
    m = outs[0]["master"][0]
    q = np.floor(np.clip(m.astype(np.float64), -1, 1) * 2147483647.0 + 0.5)   # not the tie rule; ties are measure zero here
    assert np.abs(back - (np.floor(q / 65536.0) / 32768.0)).max() <= 1.0 / 32768.0


def _pcm_case(lib, nframes, nblocks=3):
    """Three sessions, two buses each carrying a clip, rendered; returns the engine."""
    e = Engine(lib=lib, mode=K.ZL_MODE_Z, quirks=K.ZL_QUIRKS_FAITHFUL, n_sessions=3, max_frames=nframes, max_blocks=nblocks)
    rng = np.random.default_rng(5)
    for s_ in range(3):
        c = e.add_clip(rng.uniform(-1.2, 1.2, (2, 6000)).astype(np.float32), rate=48000.0)   # beyond full scale: the writer's clamps
        e.play(c, session=s_, note=60 + s_, volume=1.0)
        e.play(c, session=s_, note=67, volume=0.8, midi_channel=1)
    e.render(nblocks, nframes, e.make_clocks(nblocks, nframes))
    return e


@pytest.mark.parametrize("nframes", [256, 131])
def test_batched_pcm_read_back_is_the_bounced_files_data_chunk_oracle(oracle, nframes):
    """zl_read_master_all_pcm_async restated: row s = bytes 104.. of zl_bounce_wav(s, master) (CPU; the GPU twin is below)."""
    e = _pcm_case(oracle, nframes)
    for bits in (16, 24):
        rows = e.masters_pcm(bits)
        assert rows.shape == (3, 3 * nframes * 2 * bits // 8)
        for s_ in range(3):
            assert np.array_equal(rows[s_], e.bounce_wav(s_, -1, 0, bits)[104:104 + rows.shape[1]])
    e.close()


@pytest.mark.gpu
@pytest.mark.parametrize("nframes", [256, 131])
def test_batched_pcm_read_back_on_the_device(product, oracle, nframes):
    """Vector (frames % 4 == 0) and scalar encode paths of the batched bounce: bytes equal to the oracle's and to the
    device's own per-session zl_bounce_wav; a second render reuses the other PCM buffer."""
    e, o = _pcm_case(product, nframes), _pcm_case(oracle, nframes)
    for rep in range(3):
        for bits in (16, 24):
            rows = e.masters_pcm(bits)
            assert np.array_equal(rows, o.masters_pcm(bits))
            assert np.array_equal(rows[1], e.bounce_wav(1, -1, 0, bits)[104:104 + rows.shape[1]])
            assert len(np.unique(rows)) > 100
        for x in (e, o):
            x.render(3, nframes, x.make_clocks(3, nframes))
    e.close(); o.close()
