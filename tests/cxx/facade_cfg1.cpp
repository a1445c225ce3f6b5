// tests/cxx/facade_cfg1.cpp — a C++ host written against the reference's class names (include/zl/*.h), compiled with g++ and
// linked against libzl_b200_host.so + libzl_b200.so by tests/test_facade.py.
//   facade_cfg1 <clip.wav> <loop.wav> <out.bin>
// BASELINE configs[0] (one voice, one clip, pitch 1.0, 512-frame periods) on the global channel plus one configs[2]-style
// track (a looping clip on track 3 behind a JackPassthrough with pan, metered by AudioLevels), driven only through
// ClipAudioSource / SamplerSynth / JackPassthrough / AudioLevels / WaveFormItem.  Writes the ports, meters and a waveform mask.
// Exit codes: 0 ok, 3 = no CUDA device (the loud failure the product promises), 1 = anything else.
#include <zl/SamplerSynth.h>
#include <zl/ClipAudioSource.h>
#include <zl/JackPassthrough.h>
#include <zl/AudioLevels.h>
#include <zl/WaveFormItem.h>

#include <cstdio>
#include <cstring>
#include <vector>

static void put(FILE* f, const void* p, size_t n) { fwrite(p, 1, n, f); }

int main(int argc, char** argv) {
    if (argc < 4) return 1;
    SamplerSynth* synth = SamplerSynth::instance();
    if (!synth->initialize(nullptr, 48000.0, 512, 8)) {
        fprintf(stderr, "SamplerSynth::initialize: %s\n", synth->lastError());
        return strstr(synth->lastError(), "no CUDA device") ? 3 : 1;
    }
    ClipAudioSource clip(nullptr, nullptr, argv[1]);
    ClipAudioSource loop(nullptr, nullptr, argv[2]);
    if (clip.deviceClip() < 0 || loop.deviceClip() < 0) return 1;
    // the reference's constructor defaults, through the getters
    if (clip.slices() != 16 || clip.adsrAttack() != 0.0f || clip.adsrRelease() != 0.05f || clip.rootNote() != 60) return 1;
    // setADSRAttack resets the other three to JUCE's defaults (lib/ClipAudioSource.cpp:636-645)
    loop.setADSRAttack(0.01f);
    if (loop.adsrRelease() != 0.1f || loop.adsrDecay() != 0.1f || loop.adsrSustain() != 1.0f) return 1;
    loop.setADSRParameters(ADSRParameters(0.002f, 0.05f, 0.7f, 0.03f));
    loop.setLength(1.5f, 120);                         // 1.5 beats at 120 bpm = 0.75 s, not beat aligned
    loop.setPan(-0.4f);
    loop.setVolumeAbsolute(0.8f);
    if (loop.sliceForMidiNote(62) != 2) return 1;      // ((16 - 60 % 16) + 62) % 16
    JackPassthrough track3(3);
    track3.setDryAmount(0.8f); track3.setWetFx1Amount(0.3f); track3.setWetFx2Amount(0.0f); track3.setPanAmount(0.25f);
    if (track3.panAmount() != 0.25f || track3.muted()) return 1;

    clip.play(false);                                  // one-shot on the global un-effected channel (-2)
    loop.play(true, 3);                                // looping on track 3
    FILE* out = fopen(argv[3], "wb");
    if (!out) return 1;
    const int nblocks = 8, nframes = 512, renders = 3;
    for (int r = 0; r < renders; ++r) {
        std::vector<zl_block_clock> ck((size_t) nblocks);
        for (int b = 0; b < nblocks; ++b) {
            const uint64_t period = 10667, cur = 1000000 + (uint64_t) (r * nblocks + b) * period;
            ck[(size_t) b] = zl_block_clock{cur, cur + period, 0, 1000000, 5208};
        }
        if (r == 2) loop.stop(3);
        if (!synth->render(nblocks, nframes, ck.data())) { fprintf(stderr, "render: %s\n", synth->lastError()); return 1; }
        const size_t n = (size_t) synth->framesInLastRender();
        std::vector<float> l(n), rr(n);
        const int ports[4][2] = {{-2, ZL_BUS_RAW}, {3, ZL_BUS_RAW}, {3, ZL_BUS_DRY}, {3, ZL_BUS_FX1}};
        for (auto& p : ports) {
            if (!synth->readChannel(p[0], p[1], l.data(), rr.data())) return 1;
            put(out, l.data(), n * 4); put(out, rr.data(), n * 4);
        }
        AudioLevels::instance()->timerCallback();
        AudioLevels* a = AudioLevels::instance();
        const float lv[5] = {a->channelsA[3], a->channelsB[3], a->channels()[3], a->playbackA, a->playbackAHold};
        put(out, lv, sizeof lv);
        const float pg[2] = {loop.playbackPositionsModel()->peakGain(), (float) loop.playbackPositionsModel()->firstProgress()};
        put(out, pg, sizeof pg);
        if (synth->cpuLoad() <= 0.0f || synth->cpuLoad() > 100.0f) { fprintf(stderr, "cpuLoad %f\n", synth->cpuLoad()); return 1; }
    }
    WaveFormItem item;
    item.setSource(&loop);
    std::vector<uint8_t> mask;
    if (!item.paint(200, 64, mask)) return 1;
    put(out, mask.data(), mask.size());
    fclose(out);
    return 0;
}
