// juce_ref.cpp — thin C wrapper around the REFERENCE'S OWN JUCE objects, compiled from the sources
// where they lie under /root/reference (never copied into this repo).  Test infrastructure only:
// it pins oracle/zl_oracle.cpp's restatements (ADSR, J-mode voice, readMaxLevels, roundToInt,
// getRMSLevel) against the real reference code and generates tests/golden/ vectors.
// Built by oracle/Makefile into oracle/_ref/libzl_juceref.so (git-ignored, travels with gpurun).
#include <juce_core/juce_core.h>
#include <juce_audio_basics/juce_audio_basics.h>
#include <juce_audio_formats/juce_audio_formats.h>

#include <algorithm>
#include <cstring>
#include <memory>

namespace {

// An AudioFormatReader over caller-owned planar float PCM; everything except readSamples() is the
// reference's own base-class code (read(), readMaxLevels(), ...).
class MemoryFloatReader : public juce::AudioFormatReader {
public:
    MemoryFloatReader(const float* const* planar, int nch, juce::int64 len, double rate)
        : juce::AudioFormatReader(nullptr, "zl-memory") {
        sampleRate = rate; bitsPerSample = 32; lengthInSamples = len;
        numChannels = (unsigned int) nch; usesFloatingPointData = true;
        for (int i = 0; i < nch && i < 2; ++i) chans[i] = planar[i];
    }
    bool readSamples(int** dest, int numDestChannels, int startOffsetInDestBuffer,
                     juce::int64 startSampleInFile, int numSamples) override {
        for (int c = 0; c < numDestChannels; ++c) {
            if (dest[c] == nullptr) continue;
            float* d = reinterpret_cast<float*>(dest[c]) + startOffsetInDestBuffer;
            for (int i = 0; i < numSamples; ++i) {
                const juce::int64 s = startSampleInFile + i;
                d[i] = (c < (int) numChannels && s >= 0 && s < lengthInSamples) ? chans[c][s] : 0.0f;
            }
        }
        return true;
    }
private:
    const float* chans[2] = {nullptr, nullptr};
};

} // namespace

extern "C" {

// juce::ADSR driven like oracle's zlo_adsr_run.
int ref_adsr_run(double sr, float a, float d, float s, float r, int n_on, int n_total, int renote, float* out) {
    juce::ADSR e;
    e.setSampleRate(sr);
    e.setParameters(juce::ADSR::Parameters(a, d, s, r));
    e.noteOn();
    for (int i = 0; i < n_total; ++i) {
        if (i == n_on || (renote && i > n_on)) e.noteOff();
        out[i] = e.getNextSample();
    }
    return 0;
}

// juce::SamplerSound + juce::SamplerVoice::renderNextBlock through juce::Synthesiser.
// noteoff_block < 0: never released.  Output: nblocks*nframes per channel (stereo buffer).
int ref_jvoice_render(const float* const* planar, int nch, long long len, double rate, double outRate,
                      int root, int note, float velocity, double attack, double release,
                      int nblocks, int nframes, int noteoff_block, float* outL, float* outR) {
    MemoryFloatReader reader(planar, nch, len, rate);
    juce::BigInteger notes; notes.setRange(0, 128, true);
    juce::Synthesiser synth;
    synth.setCurrentPlaybackSampleRate(outRate);
    synth.addVoice(new juce::SamplerVoice());
    synth.addSound(new juce::SamplerSound("s", reader, notes, root, attack, release, (double) len / rate + 10.0));
    synth.noteOn(1, note, velocity);
    juce::AudioBuffer<float> buf(2, nframes);
    juce::MidiBuffer midi;
    for (int b = 0; b < nblocks; ++b) {
        if (b == noteoff_block) synth.noteOff(1, note, 0.0f, true);
        buf.clear();
        synth.renderNextBlock(buf, midi, 0, nframes);
        std::memcpy(outL + (size_t) b * nframes, buf.getReadPointer(0), sizeof(float) * (size_t) nframes);
        std::memcpy(outR + (size_t) b * nframes, buf.getReadPointer(1), sizeof(float) * (size_t) nframes);
    }
    return 0;
}

// AudioFormatReader::readMaxLevels on float PCM: out[2*c] = min, out[2*c+1] = max.
int ref_read_max_levels(const float* const* planar, int nch, long long len, double rate,
                        long long start, long long num, float* out) {
    MemoryFloatReader reader(planar, nch, len, rate);
    juce::Range<float> res[2];
    reader.readMaxLevels(start, num, res, nch);
    for (int c = 0; c < nch; ++c) { out[2 * c] = res[c].getStart(); out[2 * c + 1] = res[c].getEnd(); }
    return 0;
}

int ref_round_to_int(double v) { return juce::roundToInt(v); }
int ref_round_to_int_f(float v) { return juce::roundToInt(v); }

float ref_rms(const float* x, int n) {
    juce::AudioBuffer<float> b(1, n);
    std::memcpy(b.getWritePointer(0), x, sizeof(float) * (size_t) n);
    return b.getRMSLevel(0, 0, n);
}
float ref_magnitude(const float* x, int n) {
    juce::AudioBuffer<float> b(1, n);
    std::memcpy(b.getWritePointer(0), x, sizeof(float) * (size_t) n);
    return b.getMagnitude(0, 0, n);
}
void ref_find_min_max(const float* x, int n, float* mn, float* mx) {
    auto r = juce::FloatVectorOperations::findMinAndMax(x, n);
    *mn = r.getStart(); *mx = r.getEnd();
}

// WAV round trip through the reference's WavAudioFormat: write `bits`-bit PCM (or 32-bit float when
// bits == -32) to memory, read it back as float.  Returns frames read.
long long ref_wav_roundtrip(const float* const* planar, int nch, long long len, double rate, int bits,
                            unsigned char* wavOut, long long wavCap, long long* wavSize, float* const* back) {
    juce::WavAudioFormat fmt;
    juce::MemoryBlock block;
    {
        auto* mos = new juce::MemoryOutputStream(block, false);
        std::unique_ptr<juce::AudioFormatWriter> w(fmt.createWriterFor(mos, rate, (unsigned int) nch, bits < 0 ? 32 : bits, {}, 0));
        if (!w) { delete mos; return -1; }
        w->writeFromFloatArrays(planar, nch, (int) len);
    }
    if (wavSize) *wavSize = (long long) block.getSize();
    if (wavOut && (long long) block.getSize() <= wavCap) std::memcpy(wavOut, block.getData(), block.getSize());
    std::unique_ptr<juce::AudioFormatReader> r(fmt.createReaderFor(new juce::MemoryInputStream(block, false), true));
    if (!r) return -2;
    r->read(back, nch, 0, (int) len);
    return r->lengthInSamples;
}

// Any WAV image through the reference's reader the way SamplerSynthSound::loadSoundData does it
// (lib/SamplerSynthSound.cpp:40-46): createReaderFor, AudioBuffer<float>(min(2, numChannels), length), read(&buffer, 0,
// length, 0, true, true).  info = {numChannels, bitsPerSample, usesFloatingPointData, lengthInSamples}; back = 2 rows of
// cap floats.  Returns the frames copied out, -2 when the reader refuses the image.
long long ref_wav_read(const unsigned char* wav, long long size, long long cap, float* const* back, double* rate, long long* info) {
    juce::WavAudioFormat fmt;
    std::unique_ptr<juce::AudioFormatReader> r(fmt.createReaderFor(new juce::MemoryInputStream(wav, (size_t) size, false), true));
    if (!r) return -2;
    if (rate) *rate = r->sampleRate;
    if (info) { info[0] = r->numChannels; info[1] = r->bitsPerSample; info[2] = r->usesFloatingPointData ? 1 : 0; info[3] = r->lengthInSamples; }
    if (!(r->sampleRate > 0 && r->lengthInSamples > 0)) return 0;
    const int length = (int) r->lengthInSamples;
    juce::AudioBuffer<float> data(juce::jmin(2, (int) r->numChannels), length);
    r->read(&data, 0, length, 0, true, true);
    const long long n = std::min<long long>(cap, length);
    for (int ch = 0; ch < data.getNumChannels(); ++ch) std::memcpy(back[ch], data.getReadPointer(ch), (size_t) n * sizeof(float));
    return n;
}

} // extern "C"
