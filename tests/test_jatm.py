"""Row n4: JUCE's `jatm` thumbnail stream (AudioThumbnail::saveTo / loadFrom,
juce_AudioThumbnail.cpp:593-639).  juce_audio_utils cannot be built here, so the stream layout is checked against a
struct-level restatement of those lines in this file (parity unpinned), the CUDA path byte for byte against the oracle."""
import struct

import numpy as np
import pytest

from libzl_b200.engine import Engine, ZlError


def clip(n, nch, seed=9):
    rng = np.random.default_rng(seed)
    x = rng.uniform(-1, 1, (nch, n)).astype(np.float32)
    x[:, : n // 3] *= 0.05
    x[0, n // 2] = 1.0
    return x


def parse_jatm(b):
    """loadFrom, field by field."""
    b = bytes(b)
    assert b[:4] == b"jatm"
    spt, total, finished, n, nch, rate = struct.unpack_from("<iqqiii", b, 4)
    assert b[36:52] == bytes(16)
    data = np.frombuffer(b, np.int8, n * nch * 2, 52).reshape(n, nch, 2)
    assert len(b) == 52 + n * nch * 2
    return dict(spt=spt, total=total, finished=finished, n=n, nch=nch, rate=rate), data


CASES = [(512 * 40 + 17, 2, 512, 48000.0), (512 * 3, 1, 512, 44100.0), (100003, 2, 300, 22050.0), (511, 1, 512, 48000.0)]


@pytest.mark.parametrize("n,nch,spt,rate", CASES)
def test_oracle_stream_layout_and_round_trip(oracle, n, nch, spt, rate):
    pcm = clip(n, nch)
    e = Engine(lib=oracle)
    c = e.add_clip(pcm, rate=rate)
    e.thumb_build(c, spt)
    lv = e.thumb_level0(c)                                  # [nch, n_thumb, 2]
    hdr, data = parse_jatm(e.thumb_save_jatm(c))
    assert hdr == dict(spt=spt, total=n, finished=(n // spt) * spt, n=1 + n // spt, nch=nch, rate=int(rate))
    assert np.array_equal(data, lv.transpose(1, 0, 2))
    px = e.thumb_pixels(c, 0.0, n / rate, 257)
    # a second clip with the same audio takes the stream instead of a pass over the PCM
    c2 = e.add_clip(pcm, rate=rate)
    e.thumb_load_jatm(c2, e.thumb_save_jatm(c))
    assert np.array_equal(e.thumb_level0(c2), lv)
    assert np.array_equal(e.thumb_pixels(c2, 0.0, n / rate, 257), px)
    e.close()


def test_oracle_refuses_streams_of_other_clips(oracle):
    e = Engine(lib=oracle)
    a, b, m = e.add_clip(clip(5000, 2)), e.add_clip(clip(7000, 2)), e.add_clip(clip(5000, 1))
    s = e.thumb_save_jatm(a)                                # built on demand at 512 (lib/WaveFormItem.cpp:22)
    assert parse_jatm(s)[0]["spt"] == 512
    for target, stream in ((b, s), (m, s), (a, s[:60]), (a, np.frombuffer(b"jatx" + bytes(s[4:]), np.uint8))):
        with pytest.raises(ZlError):
            e.thumb_load_jatm(target, stream)
    e.close()


@pytest.mark.gpu
@pytest.mark.parametrize("n,nch,spt,rate", CASES + [(512 * 20000 + 5, 2, 512, 48000.0)])
def test_device_stream_equals_the_oracle_stream(product, oracle, n, nch, spt, rate):
    pcm = clip(n, nch)
    out = []
    for lib in (product, oracle):
        e = Engine(lib=lib); c = e.add_clip(pcm, rate=rate); e.thumb_build(c, spt); out.append(e.thumb_save_jatm(c)); e.close()
    assert np.array_equal(out[0], out[1])
    # load on the device what the oracle saved: same table, same pixels, no pass over the PCM
    e = Engine(lib=product)
    c = e.add_clip(pcm, rate=rate); e.thumb_build(c, spt)
    width = n // (2 * spt)                                  # zoomed out (the zoomed-in branch re-reads the PCM, :459-492)
    lv = e.thumb_level0(c)
    c2 = e.add_clip(np.zeros_like(pcm), rate=rate)          # silent audio: only the stream can produce the levels
    e.thumb_load_jatm(c2, out[1])
    assert np.array_equal(e.thumb_level0(c2), lv)
    if width >= 1:
        px = e.thumb_pixels(c, 0.0, n / rate, width)
        assert np.abs(px).max() > 1
        assert np.array_equal(e.thumb_pixels(c2, 0.0, n / rate, width), px)
    e.close()


@pytest.mark.gpu
def test_device_refuses_streams_of_other_clips(product):
    e = Engine(lib=product)
    a, b, m = e.add_clip(clip(5000, 2)), e.add_clip(clip(7000, 2)), e.add_clip(clip(5000, 1))
    s = e.thumb_save_jatm(a)
    for target, stream in ((b, s), (m, s), (a, s[:60]), (a, np.frombuffer(b"jatx" + bytes(s[4:]), np.uint8))):
        with pytest.raises(ZlError):
            e.thumb_load_jatm(target, stream)
    e.close()
