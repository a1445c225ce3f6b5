// zl_oracle.cpp — CPU ORACLE for the sampler voice-render hot path of zynthbox/libzl.
//
// THIS IS TEST INFRASTRUCTURE.  It restates, in plain C++ with no dependencies, the algorithm of
// the reference for the path named by BASELINE.json `north_star`.  Only tests/, bench.py's
// cpu_baseline / --impl reference legs and __graft_entry__.smoke() may load it; the product
// (libzl_b200/) never links, loads or calls it and has no CPU fallback.
//
// Parity pin status:
//   * juce::ADSR, juce::SamplerVoice::renderNextBlock (J-mode), readMaxLevels / findMinAndMax /
//     roundToInt / getRMSLevel are PINNED against the reference's own compiled JUCE objects
//     (oracle/ref/juce_ref.cpp -> oracle/_ref/libzl_juceref.so) and against the golden vectors
//     generated from them (tests/golden/, tests/golden/make_golden.py).
//   * libzl's own loops (SamplerSynthVoice::process, SamplerChannel, JackPassthrough, AudioLevels)
//     and juce::AudioThumbnail (juce_audio_utils) cannot be built headless here (Qt5/JACK/
//     tracktion/gui deps) and the reference ships no tests for them: "parity unpinned" for those
//     rows beyond the pinned pieces they are built from.
//
// It exports the same C ABI as the product (include/zl_b200.h) with the zlo_ prefix so a parity
// test is "same calls, two libraries, compare".  Paths below are relative to /root/reference;
// JUCE/ = tracktion_engine/modules/juce/modules/.
#define ZL_ORACLE_BUILD 1
#include "../include/zl_b200.h"

#include <algorithm>
#include <array>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <string>
#include <thread>
#include <vector>

namespace {

// ---------------------------------------------------------------------------------------------
// juce::roundToInt  JUCE/juce_core/maths/juce_MathsFunctions.h:471-485 (magic-number add).
static inline int roundToInt(double v) {
    union { int asInt[2]; double asDouble; } n;
    n.asDouble = v + 6755399441055744.0;
    return n.asInt[0];
}

// float -> uint64 conversion as on the deployment target (aarch64 fcvtzu saturates; negative -> 0).
// Used where the reference assigns a float expression to a quint64
// (lib/SamplerSynthVoice.cpp:123, :234).
static inline uint64_t f2u64(float x) {
    if (!(x > 0.0f)) return 0;
    if (x >= 18446744073709551616.0f) return UINT64_MAX;
    return (uint64_t) x;
}
static inline uint64_t d2u64(double x) {
    if (!(x > 0.0)) return 0;
    if (x >= 18446744073709551616.0) return UINT64_MAX;
    return (uint64_t) x;
}

// ---------------------------------------------------------------------------------------------
// juce::ADSR  JUCE/juce_audio_basics/utilities/juce_ADSR.h:36-267
struct Adsr {
    enum State { idle = 0, attack = 1, decay = 2, sustain = 3, release = 4 };
    State state = idle;
    float pAttack = 0.1f, pDecay = 0.1f, pSustain = 1.0f, pRelease = 0.1f;   // :66
    double sampleRate = 44100.0;                                             // :265
    float env = 0.0f, attackRate = 0.0f, decayRate = 0.0f, releaseRate = 0.0f;

    Adsr() { recalculateRates(); }                                           // :40-43

    void setParameters(float a, float d, float s, float r) {                 // :76-83
        pAttack = a; pDecay = d; pSustain = s; pRelease = r;
        recalculateRates();
    }
    bool isActive() const { return state != idle; }                          // :92
    void setSampleRate(double sr) { sampleRate = sr; }                       // :99-103 (no recalc)
    void reset() { env = 0.0f; state = idle; }                               // :107-111
    void noteOn() {                                                          // :114-129
        if (attackRate > 0.0f) state = attack;
        else if (decayRate > 0.0f) { env = 1.0f; state = decay; }
        else state = sustain;
    }
    void noteOff() {                                                         // :132-146
        if (state != idle) {
            if (pRelease > 0.0f) {
                releaseRate = (float) (env / (pRelease * sampleRate));
                state = release;
            } else {
                reset();
            }
        }
    }
    float getNextSample() {                                                  // :153-191
        if (state == idle) return 0.0f;
        if (state == attack) {
            env += attackRate;
            if (env >= 1.0f) { env = 1.0f; goToNextState(); }
        } else if (state == decay) {
            env -= decayRate;
            if (env <= pSustain) { env = pSustain; goToNextState(); }
        } else if (state == sustain) {
            env = pSustain;
        } else if (state == release) {
            env -= releaseRate;
            if (env <= 0.0f) goToNextState();
        }
        return env;
    }
    static float getRate(float distance, float timeInSeconds, double sr) {   // :232-235
        return timeInSeconds > 0.0f ? (float) (distance / (timeInSeconds * sr)) : -1.0f;
    }
    void recalculateRates() {                                                // :230-247
        attackRate = getRate(1.0f, pAttack, sampleRate);
        decayRate = getRate(1.0f - pSustain, pDecay, sampleRate);
        releaseRate = getRate(pSustain, pRelease, sampleRate);
        if ((state == attack && attackRate <= 0.0f)
            || (state == decay && (decayRate <= 0.0f || env <= pSustain))
            || (state == release && releaseRate <= 0.0f))
            goToNextState();
    }
    void goToNextState() {                                                   // :249-257
        if (state == attack) state = (decayRate > 0.0f ? decay : sustain);
        else if (state == decay) state = sustain;
        else if (state == release) reset();
    }
};

// ---------------------------------------------------------------------------------------------
// ClipAudioSource (parameter block) + SamplerSynthSound (decoded PCM)
struct Clip {
    int id = -1;
    int nch = 0;                    // min(2, file channels)     lib/SamplerSynthSound.cpp:46
    int length = 0;                 // frames                    :45
    double rate = 0.0;              // sourceSampleRate          :42
    std::vector<float> data[2];     // length + 4 zero frames of padding (juce_Sampler.cpp:46-48)
    // lib/ClipAudioSource.cpp:63-82
    float startPositionInSeconds = 0.0f;
    float lengthInSeconds = -1.0f;
    float lengthInBeats = -1.0f;
    float volumeAbsolute = 0.0f;
    float pan = 0.0f;
    int rootNote = 60;
    float adsrA = 0.0f, adsrD = 0.1f, adsrS = 1.0f, adsrR = 0.05f;           // :164-168
    std::vector<double> slicePositionsCache;
    float peakGain = 0.0f;          // ClipAudioSourcePositionsModelPrivate::peakGain

    float getDuration() const { return (float) ((double) length / rate); }   // :367 (edit length)
    float getStartPosition(int slice) const {                                // :261-268
        if (slice > -1 && slice < (int) slicePositionsCache.size())
            return (float) (startPositionInSeconds + (lengthInSeconds * slicePositionsCache[(size_t) slice]));
        return startPositionInSeconds;
    }
    float getStopPosition(int slice) const {                                 // :270-277
        if (slice > -1 && slice + 1 < (int) slicePositionsCache.size())
            return (float) (startPositionInSeconds + (lengthInSeconds * slicePositionsCache[(size_t) (slice + 1)]));
        return startPositionInSeconds + lengthInSeconds;
    }
    int startPosition(int slice) const { return (int) (getStartPosition(slice) * rate); }  // SamplerSynthSound.cpp:96-99
    int stopPosition(int slice) const { return (int) (getStopPosition(slice) * rate); }    // :101-104
    void setEvenSlices(int n) {                                              // :495-528 from empty
        slicePositionsCache.clear();
        if (n <= 0) return;
        double lastSlicePosition = 0.0f;
        double positionIncrement = (1.0f - lastSlicePosition) / (n - 0);
        double newPosition = lastSlicePosition + positionIncrement;
        slicePositionsCache.push_back(0.0f);
        while ((int) slicePositionsCache.size() < n) {
            slicePositionsCache.push_back(newPosition);
            newPosition += positionIncrement;
        }
    }
    // thumbnail level 0 (per channel (min,max) int8 pairs)
    std::vector<int8_t> thumb[2];
    int thumbSamples = 0;   // samplesPerThumbSample used
    int nThumb = 0;
};

struct Clock {
    uint64_t playhead = 0, playheadUsecs = 0, subbeatUsecs = 0;
};

static const int kMultiplier = 96;  // SyncTimer::getMultiplier  lib/SyncTimer.cpp:95,946-948

static bool equivalentTo(const zl_clip_command& a, const zl_clip_command& b) {   // lib/ClipCommand.h:33-39
    return a.clip == b.clip
        && ((a.change_slice && b.change_slice && a.slice == b.slice)
            || (!a.change_slice && !b.change_slice && a.midi_note == b.midi_note && a.midi_channel == b.midi_channel));
}

struct Engine;

// ---------------------------------------------------------------------------------------------
// SamplerSynthVoice (Z-mode)  lib/SamplerSynthVoice.cpp   /  juce::SamplerVoice (J-mode)
struct Voice {
    bool isPlaying = false;                 // lib/SamplerSynthVoice.h:32
    bool hasCommand = false;                // d->clipCommand != nullptr
    zl_clip_command cmd{};                  // *d->clipCommand
    Clip* sound = nullptr;                  // getCurrentlyPlayingSound()
    Clip* clip = nullptr;                   // d->clip
    bool hasPositionId = false;             // d->clipPositionId > -1
    uint64_t startTick = 0, nextLoopTick = 0, nextLoopUsecs = 0;
    double pitchRatio = 0, sourceSamplePosition = 0, sourceSampleLength = 0;
    float lgain = 0, rgain = 0;
    Adsr adsr;
    float posGain = 0, posProgress = 0;     // last setPositionGainAndProgress
    int midiNote = -1;

    // setCurrentCommand  lib/SamplerSynthVoice.cpp:58-98
    void setCurrentCommand(const zl_clip_command& c) {
        if (hasCommand) {
            if (c.change_looping) { cmd.looping = c.looping; cmd.change_looping = 1; }
            if (c.change_pitch) { cmd.pitch_change = c.pitch_change; cmd.change_pitch = 1; }
            if (c.change_speed) { cmd.speed_ratio = c.speed_ratio; cmd.change_speed = 1; }
            if (c.change_gain_db) { cmd.gain_db = c.gain_db; cmd.change_gain_db = 1; }
            if (c.change_volume) {
                cmd.volume = c.volume; cmd.change_volume = 1;
                lgain = cmd.volume; rgain = cmd.volume;                      // velocityToGain is identity :11-18
            }
            if (c.change_slice) cmd.slice = c.slice;
            if (c.start_playback) {
                if (sound) sourceSamplePosition = (int) (clip->getStartPosition(cmd.slice) * sound->rate);
            }
        } else {
            cmd = c; hasCommand = true;
        }
        isPlaying = hasCommand;
    }
    // Z: startNote :110-144
    void startNoteZ(int note, float velocity, Clip* s, double outRate, double pow2) {
        if (s && s->length > 0) {           // sound->isValid() && sound->clip()
            pitchRatio = pow2 * s->rate / outRate;                           // :115-116
            clip = s;
            sourceSampleLength = clip->getDuration() * s->rate;              // :120
            sourceSamplePosition = (int) (clip->getStartPosition(cmd.slice) * s->rate);   // :121
            nextLoopTick = f2u64((float) startTick + clip->lengthInBeats * (float) kMultiplier);  // :123
            nextLoopUsecs = 0;
            hasPositionId = true;                                            // :126-129
            lgain = velocity; rgain = velocity;                              // :131-132
            adsr.reset();
            adsr.setSampleRate(s->rate);
            adsr.setParameters(clip->adsrA, clip->adsrD, clip->adsrS, clip->adsrR);
            adsr.noteOn();
        }
        (void) note;
    }
    // J: juce::SamplerVoice::startNote  juce_Sampler.cpp:78-98
    void startNoteJ(float velocity, Clip* s, double outRate, double pow2) {
        pitchRatio = pow2 * s->rate / outRate;
        sourceSamplePosition = 0.0;
        lgain = velocity; rgain = velocity;
        adsr.setSampleRate(s->rate);
        adsr.setParameters(s->adsrA, s->adsrD, s->adsrS, s->adsrR);
        adsr.noteOn();
        clip = s;
        sourceSampleLength = s->getDuration() * s->rate;
    }
    // stopNote :146-169 (Z) / juce_Sampler.cpp:100-111 (J)
    void stopNote(bool allowTailOff) {
        if (allowTailOff) {
            adsr.noteOff();
        } else {
            sound = nullptr;                 // clearCurrentNote
            adsr.reset();
            clip = nullptr; hasPositionId = false;
            hasCommand = false; isPlaying = false;
            nextLoopTick = 0; nextLoopUsecs = 0;
        }
    }

    // SamplerSynthVoice::process  lib/SamplerSynthVoice.cpp:174-270.
    // L/R have nframes+1 entries when Q2 is faithful (frame f lands at index f+1).
    uint64_t processZ(float* leftBuffer, float* rightBuffer, uint32_t nframes, uint64_t current_usecs,
                      uint64_t next_usecs, const Clock& ck, int quirks) {
        uint64_t rendered = 0;
        Clip* playingSound = sound;
        if (!playingSound) return 0;
        if (!(playingSound->length > 0 && hasCommand)) return 0;
        if (nextLoopUsecs == 0) {                                            // :179-182
            const uint64_t differenceToPlayhead = nextLoopTick - ck.playhead;
            nextLoopUsecs = ck.playheadUsecs + (differenceToPlayhead * ck.subbeatUsecs);
        }
        const double microsecondsPerFrame = (double) ((next_usecs - current_usecs) / nframes);   // :183 integer division
        float peakGain = 0.0f;
        const float* const inL = playingSound->data[0].data();
        const float* const inR = playingSound->nch > 1 ? playingSound->data[1].data() : nullptr;
        const float clipVolume = clip->volumeAbsolute;                       // :189
        const int stopPosition = playingSound->stopPosition(cmd.slice);      // :190
        const int sampleDuration = playingSound->length - 1;                 // :191
        const float pan = clip->pan;
        const float lPan = (float) (0.5 * (1.0 + pan));                      // :193
        const float rPan = (float) (0.5 * (1.0 - pan));                      // :194
        const double sourceSampleRate = playingSound->rate;
        const bool isLooping = cmd.looping != 0;
        const bool q1 = (quirks & ZL_Q1_TAP_GAIN) != 0, q2 = (quirks & ZL_Q2_PREINC) != 0, q3 = (quirks & ZL_Q3_RENOTEOFF) != 0;
        bool releaseArmed = false;
        for (uint32_t frame = 0; frame < nframes; ++frame) {
            const int pos = (int) sourceSamplePosition;
            const float alpha = (float) (sourceSamplePosition - pos);
            const float invAlpha = 1.0f - alpha;
            const float envelopeValue = adsr.getNextSample();
            float l, r;
            if (q1) {                                                        // :204-205 as written
                l = sampleDuration > pos ? (inL[pos] * invAlpha + inL[pos + 1] * alpha * lgain * envelopeValue * clipVolume) : 0;
                r = (inR != nullptr && sampleDuration > pos) ? (inR[pos] * invAlpha + inR[pos + 1] * alpha * rgain * envelopeValue * clipVolume) : l;
            } else {                                                         // "fixed": gains scale the interpolated value
                l = sampleDuration > pos ? ((inL[pos] * invAlpha + inL[pos + 1] * alpha) * lgain * envelopeValue * clipVolume) : 0;
                r = (inR != nullptr && sampleDuration > pos) ? ((inR[pos] * invAlpha + inR[pos + 1] * alpha) * rgain * envelopeValue * clipVolume) : l;
            }
            const float mSignal = (float) (0.5 * (l + r));                   // :208
            const float sSignal = l - r;
            l = lPan * mSignal + sSignal;
            r = rPan * mSignal - sSignal;
            const float newGain{l + r};                                      // :213-216
            if (newGain > peakGain) peakGain = newGain;
            if (q2) { ++leftBuffer; *leftBuffer += l; ++rightBuffer; *rightBuffer += r; }   // :218-221
            else { *leftBuffer += l; ++leftBuffer; *rightBuffer += r; ++rightBuffer; }
            ++rendered;
            sourceSamplePosition += pitchRatio;                              // :223
            if (isLooping) {
                if (truncf(clip->lengthInBeats) == clip->lengthInBeats) {    // :227
                    if (current_usecs + d2u64(frame * microsecondsPerFrame) >= nextLoopUsecs) {   // :232
                        const uint64_t lengthInTicks = f2u64(clip->lengthInBeats * (float) kMultiplier);   // :234
                        nextLoopTick = nextLoopTick + lengthInTicks;
                        const uint64_t differenceToPlayhead = nextLoopTick - ck.playhead;
                        nextLoopUsecs = ck.playheadUsecs + (differenceToPlayhead * ck.subbeatUsecs);
                        sourceSamplePosition = (int) (clip->getStartPosition(cmd.slice) * sourceSampleRate);   // :241
                    }
                } else if (sourceSamplePosition >= stopPosition) {           // :243-247
                    sourceSamplePosition = (int) (clip->getStartPosition(cmd.slice) * sourceSampleRate);
                }
            } else {
                if (sourceSamplePosition >= stopPosition) {                  // :249-252
                    stopNote(false);
                    break;
                } else if (sourceSamplePosition >= (stopPosition - (adsr.pRelease * sourceSampleRate))) {   // :253-256
                    if (q3 || !releaseArmed) { stopNote(true); releaseArmed = true; }
                }
            }
            if (!adsr.isActive()) {                                          // :258-261
                stopNote(false);
                break;
            }
        }
        if (clip && hasPositionId) {                                         // :265-267
            posGain = peakGain * 0.5f;
            posProgress = (float) (sourceSamplePosition / sourceSampleLength);
        }
        return rendered;
    }

    // juce::SamplerVoice::renderNextBlock  juce_Sampler.cpp:117-163 (stereo output buffer)
    uint64_t processJ(float* outL, float* outR, int numSamples) {
        uint64_t rendered = 0;
        Clip* playingSound = sound;
        if (!playingSound) return 0;
        const float* const inL = playingSound->data[0].data();
        const float* const inR = playingSound->nch > 1 ? playingSound->data[1].data() : nullptr;
        while (--numSamples >= 0) {
            auto pos = (int) sourceSamplePosition;
            auto alpha = (float) (sourceSamplePosition - pos);
            auto invAlpha = 1.0f - alpha;
            float l = (inL[pos] * invAlpha + inL[pos + 1] * alpha);
            float r = (inR != nullptr) ? (inR[pos] * invAlpha + inR[pos + 1] * alpha) : l;
            auto envelopeValue = adsr.getNextSample();
            l *= lgain * envelopeValue;
            r *= rgain * envelopeValue;
            *outL++ += l;
            *outR++ += r;
            ++rendered;
            sourceSamplePosition += pitchRatio;
            if (sourceSamplePosition > playingSound->length) {
                stopNote(false);
                break;
            }
        }
        if (clip) { posGain = 0.0f; posProgress = (float) (sourceSamplePosition / sourceSampleLength); }
        return rendered;
    }
};

// JackPassthroughPrivate::process  lib/JackPassthrough.cpp:45-115
static void passthroughProcess(const zl_passthrough_params& p, const float* inL, const float* inR, size_t n,
                               float* dryL, float* dryR, float* fx1L, float* fx1R, float* fx2L, float* fx2R) {
    const float dryAmount = p.dry_amount, wetFx1Amount = p.wet_fx1_amount, wetFx2Amount = p.wet_fx2_amount, panAmount = p.pan_amount;
    if (p.muted) {
        memset(dryL, 0, n * 4); memset(dryR, 0, n * 4); memset(fx1L, 0, n * 4);
        memset(fx1R, 0, n * 4); memset(fx2L, 0, n * 4); memset(fx2R, 0, n * 4);
        return;
    }
    bool outputDry = true, outputWetFx1 = true, outputWetFx2 = true;
    if (panAmount == 0 && dryAmount == 0) { outputDry = false; memset(dryL, 0, n * 4); memset(dryR, 0, n * 4); }
    else if (panAmount == 0 && dryAmount == 1) { outputDry = false; memcpy(dryL, inL, n * 4); memcpy(dryR, inR, n * 4); }
    if (panAmount == 0 && wetFx1Amount == 0) { outputWetFx1 = false; memset(fx1L, 0, n * 4); memset(fx1R, 0, n * 4); }
    else if (panAmount == 0 && wetFx1Amount == 1) { outputWetFx1 = false; memcpy(fx1L, inL, n * 4); memcpy(fx1R, inR, n * 4); }
    if (panAmount == 0 && wetFx2Amount == 0) { outputWetFx2 = false; memset(fx2L, 0, n * 4); memset(fx2R, 0, n * 4); }
    else if (panAmount == 0 && wetFx2Amount == 1) { outputWetFx2 = false; memcpy(fx2L, inL, n * 4); memcpy(fx2R, inR, n * 4); }
    if (panAmount != 0 || outputDry || outputWetFx1 || outputWetFx2) {
        for (size_t frame = 0; frame < n; ++frame) {
            const float channelSampleLeft = inL[frame], channelSampleRight = inR[frame];
            if (panAmount != 0 || outputDry) {
                dryL[frame] = dryAmount * channelSampleLeft * std::min(1 - panAmount, 1.0f);
                dryR[frame] = dryAmount * channelSampleRight * std::min(1 + panAmount, 1.0f);
            }
            if (panAmount != 0 || outputWetFx1) {
                fx1L[frame] = wetFx1Amount * channelSampleLeft * std::min(1 - panAmount, 1.0f);
                fx1R[frame] = wetFx1Amount * channelSampleRight * std::min(1 + panAmount, 1.0f);
            }
            if (panAmount != 0 || outputWetFx2) {
                fx2L[frame] = wetFx2Amount * channelSampleLeft * std::min(1 - panAmount, 1.0f);
                fx2R[frame] = wetFx2Amount * channelSampleRight * std::min(1 + panAmount, 1.0f);
            }
        }
    }
}

// AudioLevels::timerCallback peak scan of one period  lib/AudioLevels.cpp:365-371
static int peakScan(const float* buf, size_t n) {
    static const float floatToIntMultiplier{131072};
    int peak = 0;
    for (size_t i = 0; i < n; ++i) {
        const int sampleValue = (int) std::abs(floatToIntMultiplier * buf[i]);
        if (sampleValue > peak) peak = sampleValue;
    }
    return peak;
}
// juce::AudioBuffer::getRMSLevel accumulation  JUCE/juce_audio_basics/buffers/juce_AudioSampleBuffer.h:1087-1095
static double sumSquares(const float* buf, size_t n) {
    double sum = 0.0;
    for (size_t i = 0; i < n; ++i) { auto sample = buf[i]; sum += sample * sample; }
    return sum;
}
static float convertTodbFS(float raw) {                                      // lib/AudioLevels.cpp:330-341
    if (raw <= 0) return -200;
    const float fValue = 20 * log10f(raw);
    if (fValue < -200) return -200;
    return fValue;
}
static float addFloat(const float& db1, const float& db2) {                  // :234-236
    return 10 * log10f(pow(10, db1 / 10) + pow(10, db2 / 10));
}

struct Meter { int peakA = 0, peakB = 0; float holdA = 0, holdB = 0; };

// SamplerChannel  lib/SamplerSynth.cpp:33-148,187-230
struct Channel {
    std::vector<Voice> voices;
    bool enabled = true;
    int midiChannel = -1;
    zl_passthrough_params pt{1.0f, 1.0f, 1.0f, 0.0f, 0};
    Meter meter;
};

struct PendingCommand { int session, bus, block; uint64_t tick; zl_clip_command c; };

} // namespace

struct zl_ctx {
    zl_config cfg{};
    std::vector<Clip*> clips;
    std::vector<Channel> channels;            // n_sessions * buses_per_session
    std::vector<PendingCommand> pending;
    std::vector<double> pow2;                 // pow(2, k/12) for k = -127..127
    std::string err;
    int threads = 1;
    // last render
    int nblocks = 0, nframes = 0;
    std::vector<float> bus[4];                // [nbus][2][n]
    std::vector<float> master;                // [nsess][2][n]
    std::vector<int32_t> peaks;               // [nblocks][nbus][2]
    std::vector<double> sumsq;
    uint64_t voiceFrames = 0, dropped = 0, totalVoiceFrames = 0;
};

namespace {

static int nbuses(const zl_ctx* c) { return c->cfg.n_sessions * c->cfg.buses_per_session; }

static Clip* findClip(zl_ctx* ctx, int id) {
    if (id < 0 || id >= (int) ctx->clips.size()) return nullptr;
    return ctx->clips[(size_t) id];
}

// SamplerChannel::handleCommand  lib/SamplerSynth.cpp:187-230
static void handleCommand(zl_ctx* ctx, Channel& ch, const zl_clip_command& c, uint64_t currentTick) {
    Clip* sound = findClip(ctx, c.clip);
    const bool jmode = ctx->cfg.mode == ZL_MODE_J;
    if (c.stop_playback || c.start_playback) {
        if (c.stop_playback) {
            if (ch.midiChannel == c.midi_channel) {
                for (Voice& v : ch.voices)
                    if (v.sound == sound && v.hasCommand && equivalentTo(v.cmd, c)) v.stopNote(true);
            }
        }
        if (c.start_playback) {
            if (ch.midiChannel == c.midi_channel) {
                for (Voice& v : ch.voices) {
                    if (!v.isPlaying) {
                        v.setCurrentCommand(c);
                        v.startTick = currentTick;
                        // juce::Synthesiser::startVoice  JUCE/juce_audio_basics/synthesisers/juce_Synthesiser.cpp:306-328
                        if (sound != nullptr) {
                            if (v.sound != nullptr) v.stopNote(false);
                            v.sound = sound;
                            v.midiNote = c.midi_note;
                            const int k = c.midi_note - sound->rootNote;
                            const double p2 = ctx->pow2[(size_t) (std::min(127, std::max(-127, k)) + 127)];
                            if (jmode) v.startNoteJ(c.volume, sound, ctx->cfg.sample_rate, p2);
                            else v.startNoteZ(c.midi_note, c.volume, sound, ctx->cfg.sample_rate, p2);
                        }
                        break;
                    }
                }
            }
        }
    } else {
        if (ch.midiChannel == c.midi_channel) {
            for (Voice& v : ch.voices)
                if (v.sound == sound && v.hasCommand && equivalentTo(v.cmd, c)) v.setCurrentCommand(c);
        }
    }
}

static void renderSession(zl_ctx* ctx, int s, int nblocks, int nframes, const zl_block_clock* clocks,
                          const std::vector<std::vector<const PendingCommand*>>& cmdsByBus, uint64_t* vfOut) {
    const int B = ctx->cfg.buses_per_session;
    const size_t n = (size_t) nblocks * (size_t) nframes;
    const int q = ctx->cfg.mode == ZL_MODE_Z ? ctx->cfg.quirks : 0;
    const bool q2 = (q & ZL_Q2_PREINC) != 0;
    std::vector<float> L((size_t) nframes + 1), R((size_t) nframes + 1);
    std::vector<float> tmp[6];
    for (auto& t : tmp) t.resize((size_t) nframes);
    std::vector<float> mL((size_t) nframes), mR((size_t) nframes);
    uint64_t vf = 0;
    std::vector<size_t> cursor((size_t) B, 0);
    for (int b = 0; b < nblocks; ++b) {
        const zl_block_clock& ck = clocks[b];
        Clock clock{ck.playhead, ck.playhead_usecs, ck.subbeat_usecs};
        for (int bi = 0; bi < B; ++bi) {
            const int gb = s * B + bi;
            Channel& ch = ctx->channels[(size_t) gb];
            // drain the command ring  lib/SamplerSynth.cpp:118-122
            const auto& list = cmdsByBus[(size_t) gb];
            size_t& cur = cursor[(size_t) bi];
            while (cur < list.size() && list[cur]->block == b) {
                handleCommand(ctx, ch, list[cur]->c, list[cur]->tick);
                ++cur;
            }
            std::fill(L.begin(), L.end(), 0.0f);                             // :134-135
            std::fill(R.begin(), R.end(), 0.0f);
            if (ch.enabled) {
                for (Voice& v : ch.voices) {                                 // :136-140
                    if (v.isPlaying) {
                        if (ctx->cfg.mode == ZL_MODE_J) vf += v.processJ(L.data(), R.data(), nframes);
                        else vf += v.processZ(L.data(), R.data(), (uint32_t) nframes, ck.current_usecs, ck.next_usecs, clock, q);
                    }
                }
            }
            (void) q2;  // with Q2 frame f sits at index f+1; the JACK buffer is indices [0, nframes)
            float* o;
            const size_t off = (size_t) b * (size_t) nframes;
            if (ctx->cfg.outputs & ZL_OUT_RAW) {
                o = ctx->bus[0].data() + ((size_t) gb * 2) * n + off; memcpy(o, L.data(), (size_t) nframes * 4);
                o = ctx->bus[0].data() + ((size_t) gb * 2 + 1) * n + off; memcpy(o, R.data(), (size_t) nframes * 4);
            }
            passthroughProcess(ch.pt, L.data(), R.data(), (size_t) nframes, tmp[0].data(), tmp[1].data(),
                               tmp[2].data(), tmp[3].data(), tmp[4].data(), tmp[5].data());
            for (int w = 0; w < 3; ++w) {
                if (ctx->cfg.outputs & (ZL_OUT_DRY << w)) {
                    o = ctx->bus[1 + w].data() + ((size_t) gb * 2) * n + off; memcpy(o, tmp[2 * w].data(), (size_t) nframes * 4);
                    o = ctx->bus[1 + w].data() + ((size_t) gb * 2 + 1) * n + off; memcpy(o, tmp[2 * w + 1].data(), (size_t) nframes * 4);
                }
            }
            // JACK port mix: first connection copied, the others added, in bus order.
            if (bi == 0) { mL = tmp[0]; mR = tmp[1]; }
            else for (int f = 0; f < nframes; ++f) { mL[(size_t) f] += tmp[0][(size_t) f]; mR[(size_t) f] += tmp[1][(size_t) f]; }
            if (ctx->cfg.outputs & ZL_OUT_LEVELS) {
                int32_t* pk = ctx->peaks.data() + ((size_t) b * (size_t) nbuses(ctx) + (size_t) gb) * 2;
                pk[0] = peakScan(tmp[0].data(), (size_t) nframes);
                pk[1] = peakScan(tmp[1].data(), (size_t) nframes);
            }
            if (ctx->cfg.outputs & ZL_OUT_RMS) {
                double* sq = ctx->sumsq.data() + ((size_t) b * (size_t) nbuses(ctx) + (size_t) gb) * 2;
                sq[0] = sumSquares(tmp[0].data(), (size_t) nframes);
                sq[1] = sumSquares(tmp[1].data(), (size_t) nframes);
            }
        }
        if (ctx->cfg.outputs & ZL_OUT_MASTER) {
            const size_t off = (size_t) b * (size_t) nframes;
            memcpy(ctx->master.data() + ((size_t) s * 2) * n + off, mL.data(), (size_t) nframes * 4);
            memcpy(ctx->master.data() + ((size_t) s * 2 + 1) * n + off, mR.data(), (size_t) nframes * 4);
        }
    }
    *vfOut = vf;
}

// juce::AudioThumbnail::MinMaxValue::setFloat  JUCE/juce_audio_utils/gui/juce_AudioThumbnail.cpp:46-66
static void setFloat(float mn, float mx, int8_t* out) {
    int a = roundToInt(mn * 127.0f), b = roundToInt(mx * 127.0f);
    a = std::max(-128, std::min(127, a)); b = std::max(-128, std::min(127, b));
    int8_t v0 = (int8_t) a, v1 = (int8_t) b;
    if (v0 == v1) { if (v1 == 127) v0--; else v1++; }
    out[0] = v0; out[1] = v1;
}
// AudioFormatReader::readMaxLevels (float data; samples past the end read as 0)
// JUCE/juce_audio_formats/format/juce_AudioFormatReader.cpp:199-249
static void readMaxLevels(const Clip* c, int ch, int64_t start, int64_t num, float* mn, float* mx) {
    if (num <= 0) { *mn = 0; *mx = 0; return; }
    float lo = 0, hi = 0; bool first = true;
    for (int64_t i = start; i < start + num; ++i) {
        const float v = (i >= 0 && i < c->length) ? c->data[ch][(size_t) i] : 0.0f;
        if (first) { lo = hi = v; first = false; }
        else { if (v < lo) lo = v; if (v > hi) hi = v; }
    }
    *mn = lo; *mx = hi;
}
// LevelDataSource::readNextBlock + createChannels sizing  juce_AudioThumbnail.cpp:225-269,666-670
static void buildThumb(Clip* c, int spt) {
    c->thumbSamples = spt;
    c->nThumb = 1 + (int) (c->length / spt);
    const int done = c->length / spt;     // thumb samples actually computed; the last stays (0,0)
    for (int ch = 0; ch < c->nch; ++ch) {
        c->thumb[ch].assign((size_t) c->nThumb * 2, 0);
        for (int i = 0; i < done; ++i) {
            float mn, mx;
            readMaxLevels(c, ch, (int64_t) i * spt, spt, &mn, &mx);
            setFloat(mn, mx, &c->thumb[ch][(size_t) i * 2]);
        }
    }
}
// ThumbData::getMinMax  juce_AudioThumbnail.cpp:292-319
static void getMinMax(const Clip* c, int ch, int startSample, int endSample, int8_t* out) {
    if (startSample >= 0) {
        endSample = std::min(endSample, c->nThumb - 1);
        int8_t mx = -128, mn = 127;
        while (startSample <= endSample) {
            const int8_t lo = c->thumb[ch][(size_t) startSample * 2], hi = c->thumb[ch][(size_t) startSample * 2 + 1];
            if (lo < mn) mn = lo;
            if (hi > mx) mx = hi;
            ++startSample;
        }
        if (mn <= mx) { out[0] = mn; out[1] = mx; return; }
    }
    out[0] = 1; out[1] = 0;
}
// CachedWindow::refillCache  juce_AudioThumbnail.cpp:430-521
static bool refillCache(Clip* c, int numSamples, double startTime, double endTime, int8_t* out) {
    const double rate = c->rate;
    const int sampsPerThumbSample = c->thumbSamples;
    auto timePerPixel = (endTime - startTime) / numSamples;
    if (numSamples <= 0 || timePerPixel <= 0.0 || rate <= 0) return false;
    memset(out, 0, (size_t) c->nch * (size_t) numSamples * 2);
    if (timePerPixel * rate <= sampsPerThumbSample) {                        // :459-492 (levelData present)
        auto sample = roundToInt(startTime * rate);
        for (int i = 0; i < numSamples; ++i) {
            auto nextSample = roundToInt((startTime + timePerPixel) * rate);
            if (sample >= 0) {
                if (sample >= c->length) {
                    for (int chan = 0; chan < c->nch; ++chan) { out[((size_t) chan * numSamples + i) * 2] = 0; out[((size_t) chan * numSamples + i) * 2 + 1] = 0; }
                } else {
                    for (int chan = 0; chan < c->nch; ++chan) {
                        float mn, mx;
                        readMaxLevels(c, chan, sample, std::max(1, nextSample - sample), &mn, &mx);
                        setFloat(mn, mx, &out[((size_t) chan * numSamples + i) * 2]);
                    }
                }
            }
            startTime += timePerPixel;
            sample = nextSample;
        }
    } else {                                                                 // :493-518
        const double cachedStart = startTime;
        for (int channelNum = 0; channelNum < c->nch; ++channelNum) {
            auto timeToThumbSampleFactor = rate / (double) sampsPerThumbSample;
            startTime = cachedStart;
            auto sample = roundToInt(startTime * timeToThumbSampleFactor);
            int8_t* cacheData = out + (size_t) channelNum * numSamples * 2;
            for (int i = numSamples; --i >= 0;) {
                auto nextSample = roundToInt((startTime + timePerPixel) * timeToThumbSampleFactor);
                getMinMax(c, channelNum, sample, nextSample, cacheData);
                cacheData += 2;
                startTime += timePerPixel;
                sample = nextSample;
            }
        }
    }
    return true;
}

} // namespace

// =============================================================================================
// C ABI (zlo_*)
extern "C" {

int zlo_config_default(zl_config* c) {
    if (!c) return ZL_ERR_INVALID;
    memset(c, 0, sizeof(*c));
    c->sample_rate = 48000.0; c->n_sessions = 1; c->buses_per_session = 12; c->voices_per_bus = 8;
    c->max_clips = 256; c->max_slices = 16; c->max_frames = 1024; c->max_blocks = 64;
    c->mode = ZL_MODE_Z; c->quirks = ZL_QUIRKS_FAITHFUL; c->outputs = ZL_OUT_ALL; c->device = 0;
    return ZL_OK;
}

int zlo_ctx_create(const zl_config* cfg, zl_ctx** out) {
    if (!cfg || !out) return ZL_ERR_INVALID;
    const int V = cfg->voices_per_bus;
    if (cfg->n_sessions < 1 || cfg->buses_per_session < 1 || V < 1 || V > 1024 || cfg->sample_rate <= 0) return ZL_ERR_INVALID;
    zl_ctx* c = new zl_ctx;
    c->cfg = *cfg;
    c->channels.resize((size_t) cfg->n_sessions * (size_t) cfg->buses_per_session);
    for (size_t i = 0; i < c->channels.size(); ++i) {
        c->channels[i].voices.resize((size_t) V);
        c->channels[i].midiChannel = (int) (i % (size_t) cfg->buses_per_session) - 2;      // lib/SamplerSynth.cpp:270
    }
    c->pow2.resize(255);
    for (int k = -127; k <= 127; ++k) c->pow2[(size_t) (k + 127)] = std::pow(2.0, k / 12.0);  // lib/SamplerSynthVoice.cpp:115
    *out = c;
    return ZL_OK;
}
int zlo_ctx_destroy(zl_ctx* c) { if (!c) return ZL_ERR_INVALID; for (Clip* k : c->clips) delete k; delete c; return ZL_OK; }
const char* zlo_last_error(zl_ctx* c) { return c ? c->err.c_str() : "null ctx"; }
int zlo_ctx_set_stream(zl_ctx*, void*) { return ZL_OK; }
int zlo_ctx_set_overlap(zl_ctx*, int32_t) { return ZL_OK; }
int zlo_ctx_set_graph(zl_ctx*, int32_t) { return ZL_OK; }
int zlo_sync(zl_ctx*) { return ZL_OK; }
ZL_EXPORT int zlo_set_threads(zl_ctx* c, int n) { if (!c || n < 1) return ZL_ERR_INVALID; c->threads = n; return ZL_OK; }

int zlo_clip_create_from_pcm(zl_ctx* ctx, const float* const* planar, int32_t nch, int64_t length, double rate, int32_t* clip_id) {
    if (!ctx || !planar || nch < 1 || length < 1 || length > 2147483647LL - 8 || rate <= 0 || !clip_id) return ZL_ERR_INVALID;
    if ((int) ctx->clips.size() >= ctx->cfg.max_clips) return ZL_ERR_STATE;
    Clip* c = new Clip;
    c->nch = std::min(2, nch); c->length = (int) length; c->rate = rate;
    for (int ch = 0; ch < c->nch; ++ch) {
        c->data[ch].assign((size_t) length + 4, 0.0f);
        memcpy(c->data[ch].data(), planar[ch], (size_t) length * 4);
    }
    c->lengthInSeconds = c->getDuration();                                   // lib/ClipAudioSource.cpp:158
    c->volumeAbsolute = std::exp((-3.0f - 6.0f) * (1.0f / 20.0f));           // tracktion default master fader (-3 dB): tracktion_Edit.cpp:826, tracktion_AudioUtilities.cpp:138-141
    c->setEvenSlices(std::min(16, ctx->cfg.max_slices));                     // :204
    c->id = (int) ctx->clips.size();
    ctx->clips.push_back(c);
    *clip_id = c->id;
    return ZL_OK;
}
int zlo_clip_create_from_device(zl_ctx*, const void*, int32_t, int64_t, double, int32_t*) { return ZL_ERR_INVALID; }
int zlo_clip_destroy(zl_ctx* ctx, int32_t id) {
    Clip* c = ctx ? findClip(ctx, id) : nullptr; if (!c) return ZL_ERR_INVALID;
    // ~ClipAudioSource calls stop() and unregisterClip (lib/ClipAudioSource.cpp:207-210), after which a voice still holding the
    // clip would read freed memory; the boundary defines it instead: voices still on the clip end as by stopNote(0, false).
    for (Channel& ch : ctx->channels)
        for (Voice& v : ch.voices)
            if (v.sound == c || v.clip == c) v.stopNote(false);
    delete c; ctx->clips[(size_t) id] = nullptr; return ZL_OK;
}
static void fillParams(const Clip* c, zl_clip_params* p) {
    memset(p, 0, sizeof(*p));
    p->start_s = c->startPositionInSeconds; p->length_s = c->lengthInSeconds; p->length_beats = c->lengthInBeats;
    p->pan = c->pan; p->volume_abs = c->volumeAbsolute; p->attack = c->adsrA; p->decay = c->adsrD; p->sustain = c->adsrS;
    p->release = c->adsrR; p->root_note = c->rootNote; p->n_slices = (int) c->slicePositionsCache.size(); p->slice_pos = nullptr;
}
int zlo_clip_params_default(zl_ctx* ctx, int32_t id, zl_clip_params* p) {
    Clip* c = ctx ? findClip(ctx, id) : nullptr; if (!c || !p) return ZL_ERR_INVALID;
    Clip d; d.length = c->length; d.rate = c->rate; d.lengthInSeconds = d.getDuration();
    d.volumeAbsolute = std::exp((-3.0f - 6.0f) * (1.0f / 20.0f)); d.setEvenSlices(std::min(16, ctx->cfg.max_slices));
    fillParams(&d, p); return ZL_OK;
}
int zlo_clip_get_params(zl_ctx* ctx, int32_t id, zl_clip_params* p) {
    Clip* c = ctx ? findClip(ctx, id) : nullptr; if (!c || !p) return ZL_ERR_INVALID;
    fillParams(c, p); return ZL_OK;
}
int zlo_clip_set_params(zl_ctx* ctx, int32_t id, const zl_clip_params* p) {
    Clip* c = ctx ? findClip(ctx, id) : nullptr; if (!c || !p) return ZL_ERR_INVALID;
    if (p->n_slices < 0 || p->n_slices > ctx->cfg.max_slices) return ZL_ERR_INVALID;
    c->startPositionInSeconds = std::max(0.0f, p->start_s);                  // :256
    c->lengthInSeconds = p->length_s < 0 ? c->getDuration() : p->length_s;
    c->lengthInBeats = p->length_beats; c->pan = p->pan; c->volumeAbsolute = p->volume_abs;
    c->adsrA = p->attack; c->adsrD = p->decay; c->adsrS = p->sustain; c->adsrR = p->release; c->rootNote = p->root_note;
    if (p->slice_pos) c->slicePositionsCache.assign(p->slice_pos, p->slice_pos + p->n_slices);
    else c->setEvenSlices(p->n_slices);
    return ZL_OK;
}
int zlo_clip_slice_bounds(zl_ctx* ctx, int32_t id, int32_t slice, int32_t* a, int32_t* b) {
    Clip* c = ctx ? findClip(ctx, id) : nullptr; if (!c) return ZL_ERR_INVALID;
    if (a) *a = c->startPosition(slice);
    if (b) *b = c->stopPosition(slice);
    return ZL_OK;
}

int zlo_clip_command_default(zl_clip_command* c) {                           // lib/ClipCommand.h:14-31
    if (!c) return ZL_ERR_INVALID;
    memset(c, 0, sizeof(*c)); c->clip = -1; c->midi_note = -1; c->midi_channel = -1; c->slice = -1; return ZL_OK;
}
int zlo_command(zl_ctx* ctx, int32_t session, int32_t block, uint64_t tick, const zl_clip_command* cmd) {
    if (!ctx || !cmd || session < 0 || session >= ctx->cfg.n_sessions || block < 0) return ZL_ERR_INVALID;
    // SamplerSynth::handleClipCommand  lib/SamplerSynth.cpp:328-341
    const int bus = cmd->midi_channel + 2;
    if (!findClip(ctx, cmd->clip) || bus < 0 || bus >= ctx->cfg.buses_per_session) { ctx->dropped++; return ZL_OK; }
    ctx->pending.push_back(PendingCommand{session, bus, block, tick, *cmd});
    return ZL_OK;
}
int zlo_command_batch(zl_ctx* ctx, int32_t n, const int32_t* s, const int32_t* b, const uint64_t* t, const zl_clip_command* c) {
    if (!ctx || n < 0 || (n && (!s || !b || !t || !c))) return ZL_ERR_INVALID;
    for (int i = 0; i < n; ++i) { int rc = zlo_command(ctx, s[i], b[i], t[i], &c[i]); if (rc != ZL_OK) return rc; }
    return ZL_OK;
}
int zlo_total_counters(zl_ctx* ctx, uint64_t* vf, uint64_t* sf) { if (!ctx) return ZL_ERR_INVALID; if (vf) *vf = ctx->totalVoiceFrames; if (sf) *sf = 0; return ZL_OK; }
int zlo_last_source_floats(zl_ctx*, uint64_t* n) { if (!n) return ZL_ERR_INVALID; *n = 0; return ZL_OK; }
int zlo_bus_set_enabled(zl_ctx* ctx, int32_t s, int32_t b, int32_t en) {
    if (!ctx || s < 0 || s >= ctx->cfg.n_sessions || b < 0 || b >= ctx->cfg.buses_per_session) return ZL_ERR_INVALID;
    ctx->channels[(size_t) s * (size_t) ctx->cfg.buses_per_session + (size_t) b].enabled = en != 0; return ZL_OK;
}
int zlo_passthrough_set(zl_ctx* ctx, int32_t s, int32_t b, const zl_passthrough_params* p) {
    if (!ctx || !p || s < 0 || s >= ctx->cfg.n_sessions || b < 0 || b >= ctx->cfg.buses_per_session) return ZL_ERR_INVALID;
    ctx->channels[(size_t) s * (size_t) ctx->cfg.buses_per_session + (size_t) b].pt = *p; return ZL_OK;
}
int zlo_passthrough_get(zl_ctx* ctx, int32_t s, int32_t b, zl_passthrough_params* p) {
    if (!ctx || !p || s < 0 || s >= ctx->cfg.n_sessions || b < 0 || b >= ctx->cfg.buses_per_session) return ZL_ERR_INVALID;
    *p = ctx->channels[(size_t) s * (size_t) ctx->cfg.buses_per_session + (size_t) b].pt; return ZL_OK;
}

int zlo_render(zl_ctx* ctx, int32_t nblocks, int32_t nframes, const zl_block_clock* clocks) {
    if (!ctx || nblocks < 1 || nframes < 1 || !clocks || nblocks > ctx->cfg.max_blocks || nframes > ctx->cfg.max_frames) return ZL_ERR_INVALID;
    const size_t n = (size_t) nblocks * (size_t) nframes;
    const size_t NB = (size_t) nbuses(ctx);
    ctx->nblocks = nblocks; ctx->nframes = nframes;
    for (int w = 0; w < 4; ++w)
        if (ctx->cfg.outputs & (ZL_OUT_RAW << w)) ctx->bus[w].assign(NB * 2 * n, 0.0f);
    if (ctx->cfg.outputs & ZL_OUT_MASTER) ctx->master.assign((size_t) ctx->cfg.n_sessions * 2 * n, 0.0f);
    ctx->peaks.assign((size_t) nblocks * NB * 2, 0);
    ctx->sumsq.assign((size_t) nblocks * NB * 2, 0.0);
    // per-bus command lists, stable by arrival, then by block (ring capacity 256: lib/SamplerSynth.cpp:32,332-334)
    std::vector<std::vector<const PendingCommand*>> byBus(NB);
    std::stable_sort(ctx->pending.begin(), ctx->pending.end(), [](const PendingCommand& a, const PendingCommand& b) { return a.block < b.block; });
    {
        std::vector<int> countInBlock(NB, 0), lastBlock(NB, -1);
        for (const PendingCommand& p : ctx->pending) {
            const size_t gb = (size_t) p.session * (size_t) ctx->cfg.buses_per_session + (size_t) p.bus;
            if (p.block >= nblocks) { ctx->dropped++; continue; }
            if (lastBlock[gb] != p.block) { lastBlock[gb] = p.block; countInBlock[gb] = 0; }
            if (++countInBlock[gb] > 256) { ctx->dropped++; continue; }
            byBus[gb].push_back(&p);
        }
    }
    const int S = ctx->cfg.n_sessions;
    std::vector<uint64_t> vf((size_t) S, 0);
    const int T = std::max(1, std::min(ctx->threads, S));
    if (T == 1) {
        for (int s = 0; s < S; ++s) renderSession(ctx, s, nblocks, nframes, clocks, byBus, &vf[(size_t) s]);
    } else {
        std::atomic<int> next{0};
        std::vector<std::thread> pool;
        for (int t = 0; t < T; ++t)
            pool.emplace_back([&]() { for (;;) { int s = next.fetch_add(1); if (s >= S) break; renderSession(ctx, s, nblocks, nframes, clocks, byBus, &vf[(size_t) s]); } });
        for (auto& th : pool) th.join();
    }
    ctx->voiceFrames = 0;
    for (uint64_t v : vf) ctx->voiceFrames += v;
    ctx->totalVoiceFrames += ctx->voiceFrames;
    ctx->pending.clear();
    return ZL_OK;
}

int zlo_read_bus(zl_ctx* ctx, int32_t s, int32_t b, int32_t which, float* l, float* r, int64_t n) {
    if (!ctx || which < 0 || which > 3 || s < 0 || s >= ctx->cfg.n_sessions || b < 0 || b >= ctx->cfg.buses_per_session) return ZL_ERR_INVALID;
    const size_t N = (size_t) ctx->nblocks * (size_t) ctx->nframes;
    if ((size_t) n != N || !(ctx->cfg.outputs & (ZL_OUT_RAW << which))) return ZL_ERR_STATE;
    const size_t gb = (size_t) s * (size_t) ctx->cfg.buses_per_session + (size_t) b;
    if (l) memcpy(l, ctx->bus[which].data() + gb * 2 * N, N * 4);
    if (r) memcpy(r, ctx->bus[which].data() + (gb * 2 + 1) * N, N * 4);
    return ZL_OK;
}
int zlo_read_master(zl_ctx* ctx, int32_t s, float* l, float* r, int64_t n) {
    if (!ctx || s < 0 || s >= ctx->cfg.n_sessions) return ZL_ERR_INVALID;
    const size_t N = (size_t) ctx->nblocks * (size_t) ctx->nframes;
    if ((size_t) n != N || !(ctx->cfg.outputs & ZL_OUT_MASTER)) return ZL_ERR_STATE;
    if (l) memcpy(l, ctx->master.data() + (size_t) s * 2 * N, N * 4);
    if (r) memcpy(r, ctx->master.data() + ((size_t) s * 2 + 1) * N, N * 4);
    return ZL_OK;
}
int zlo_read_master_all(zl_ctx* ctx, float* out, int64_t n) {
    if (!ctx || !out) return ZL_ERR_INVALID;
    const size_t N = (size_t) ctx->nblocks * (size_t) ctx->nframes;
    if ((size_t) n != N || !(ctx->cfg.outputs & ZL_OUT_MASTER)) return ZL_ERR_STATE;
    memcpy(out, ctx->master.data(), ctx->master.size() * 4); return ZL_OK;
}
int zlo_read_master_all_async(zl_ctx* ctx, float* out, int64_t n) { return zlo_read_master_all(ctx, out, n); }   // nothing to overlap on the CPU
int zlo_bounce_wav(zl_ctx* ctx, int32_t s, int32_t b, int32_t which, int32_t bits, void* out, uint64_t cap, uint64_t* size);
// every session's master as the data chunk the reference's writer would produce for it: zlo_bounce_wav's bytes without the header
int zlo_read_master_all_pcm_async(zl_ctx* ctx, void* out, int64_t n, int32_t bits) {
    if (!ctx || !out || (bits != 16 && bits != 24)) return ZL_ERR_INVALID;
    const size_t N = (size_t) ctx->nblocks * (size_t) ctx->nframes;
    if ((size_t) n != N || !(ctx->cfg.outputs & ZL_OUT_MASTER)) return ZL_ERR_STATE;
    const size_t body = N * 2 * (size_t) (bits / 8);
    std::vector<unsigned char> file(104 + body + 1);
    for (int s = 0; s < ctx->cfg.n_sessions; ++s) {
        uint64_t size = 0;
        const int rc = zlo_bounce_wav(ctx, s, -1, 0, bits, file.data(), file.size(), &size);
        if (rc != ZL_OK) return rc;
        memcpy(static_cast<unsigned char*>(out) + (size_t) s * body, file.data() + 104, body);
    }
    return ZL_OK;
}
int zlo_read_wait(zl_ctx* ctx) { return ctx ? ZL_OK : ZL_ERR_INVALID; }
int zlo_device_bus_ptr(zl_ctx*, int32_t, float**, int64_t*) { return ZL_ERR_INVALID; }
int zlo_device_master_ptr(zl_ctx*, float**, int64_t*) { return ZL_ERR_INVALID; }
int zlo_read_block_levels(zl_ctx* ctx, int32_t* peaks, double* sumsq) {
    if (!ctx) return ZL_ERR_INVALID;
    if (peaks) memcpy(peaks, ctx->peaks.data(), ctx->peaks.size() * 4);
    if (sumsq) memcpy(sumsq, ctx->sumsq.data(), ctx->sumsq.size() * 8);
    return ZL_OK;
}
int zlo_levels_tick(zl_ctx* ctx, int32_t s, int32_t b, int32_t block, int32_t isPlayback, zl_levels* out) {
    if (!ctx || !out || s < 0 || s >= ctx->cfg.n_sessions || b < 0 || b >= ctx->cfg.buses_per_session || block >= ctx->nblocks) return ZL_ERR_INVALID;
    static const float intToFloatMultiplier{0.00000152587};                  // lib/AudioLevels.cpp:349
    const size_t gb = (size_t) s * (size_t) ctx->cfg.buses_per_session + (size_t) b;
    Meter& m = ctx->channels[gb].meter;
    m.peakA = std::max(0, m.peakA - 10000);                                  // :359-360
    m.peakB = std::max(0, m.peakB - 10000);
    float rmsA = 0, rmsB = 0;
    if (block >= 0) {
        const int32_t* pk = ctx->peaks.data() + ((size_t) block * (size_t) nbuses(ctx) + gb) * 2;
        if (pk[0] > m.peakA) m.peakA = pk[0];
        if (pk[1] > m.peakB) m.peakB = pk[1];
        const double* sq = ctx->sumsq.data() + ((size_t) block * (size_t) nbuses(ctx) + gb) * 2;
        rmsA = (float) std::sqrt(sq[0] / ctx->nframes); rmsB = (float) std::sqrt(sq[1] / ctx->nframes);
    }
    const float peakA{m.peakA * intToFloatMultiplier}, peakB{m.peakB * intToFloatMultiplier};   // :385
    const float peakDbA{convertTodbFS(peakA)}, peakDbB{convertTodbFS(peakB)};
    memset(out, 0, sizeof(*out));
    out->peak_a = m.peakA; out->peak_b = m.peakB;
    out->db_a = peakDbA <= -200 ? -200 : peakDbA; out->db_b = peakDbB <= -200 ? -200 : peakDbB;
    out->db_sum = isPlayback ? addFloat(peakDbA, peakDbB) : addFloat(out->db_a, out->db_b);   // :394 / :406
    if (isPlayback) {                                                        // :395-396
        m.holdA = (peakA >= m.holdA) ? peakA : m.holdA * 0.9f;
        m.holdB = (peakB >= m.holdB) ? peakB : m.holdB * 0.9f;
    }
    out->hold_a = m.holdA; out->hold_b = m.holdB; out->rms_a = rmsA; out->rms_b = rmsB;
    return ZL_OK;
}
int zlo_read_voices(zl_ctx* ctx, zl_voice_info* out, int32_t cap) {
    if (!ctx || !out) return ZL_ERR_INVALID;
    int i = 0;
    for (Channel& ch : ctx->channels)
        for (Voice& v : ch.voices) {
            if (i >= cap) return ZL_ERR_INVALID;
            zl_voice_info& o = out[i++];
            memset(&o, 0, sizeof(o));
            o.position = v.sourceSamplePosition; o.pitch_ratio = v.pitchRatio; o.envelope = v.adsr.env; o.gain = v.lgain;
            o.pos_gain = v.posGain; o.pos_progress = v.posProgress; o.playing = v.isPlaying ? 1 : 0;
            o.clip = v.sound ? v.sound->id : -1; o.adsr_state = (int) v.adsr.state; o.midi_note = v.midiNote;
            o.slice = v.cmd.slice; o.looping = v.cmd.looping;
        }
    return ZL_OK;
}
int zlo_read_positions(zl_ctx* ctx, int32_t clip, float* gains, float* progress, int32_t cap, int32_t* count) {
    if (!ctx || !count) return ZL_ERR_INVALID;
    int n = 0;
    for (Channel& ch : ctx->channels)
        for (Voice& v : ch.voices)
            if (v.isPlaying && v.clip && v.clip->id == clip && v.hasPositionId && n < cap && n < 32) {   // POSITION_COUNT 32
                if (gains) gains[n] = v.posGain;
                if (progress) progress[n] = v.posProgress;
                ++n;
            }
    *count = n; return ZL_OK;
}
// ClipAudioSourcePositionsModel::peakGain / firstProgress   lib/ClipAudioSourcePositionsModel.cpp:160-185
int zlo_clip_peak_gain(zl_ctx* ctx, int32_t clip, float* peakOut, double* firstProgress) {
    Clip* c = ctx ? findClip(ctx, clip) : nullptr; if (!c || !peakOut) return ZL_ERR_INVALID;
    float peak{0.0f};
    double progress{-1.0f};
    int n = 0;
    for (Channel& ch : ctx->channels)
        for (Voice& v : ch.voices)
            if (v.isPlaying && v.clip == c && v.hasPositionId && n < 32) {
                peak = std::max(peak, v.posGain);
                if (n == 0) progress = v.posProgress;
                ++n;
            }
    if (std::abs(c->peakGain - peak) > 0.01) c->peakGain = peak;
    *peakOut = c->peakGain;
    if (firstProgress) *firstProgress = progress;
    return ZL_OK;
}
int zlo_kernel_timing(zl_ctx* ctx, int32_t) { return ctx ? ZL_OK : ZL_ERR_INVALID; }   // no kernels in the oracle
int zlo_kernel_time(zl_ctx* ctx, double* ms, uint64_t* n) { if (!ctx) return ZL_ERR_INVALID; if (ms) *ms = 0.0; if (n) *n = 0; return ZL_OK; }
int zlo_last_voice_frames(zl_ctx* ctx, uint64_t* vf) { if (!ctx || !vf) return ZL_ERR_INVALID; *vf = ctx->voiceFrames; return ZL_OK; }
int zlo_dropped_commands(zl_ctx* ctx, uint64_t* n) { if (!ctx || !n) return ZL_ERR_INVALID; *n = ctx->dropped; return ZL_OK; }
int zlo_launch_count(zl_ctx*, uint64_t* n) { if (n) *n = 0; return ZL_OK; }
int zlo_host_profile(zl_ctx* ctx, double* a, double* b, uint64_t* n) { if (!ctx) return ZL_ERR_INVALID; if (a) *a = 0; if (b) *b = 0; if (n) *n = 0; return ZL_OK; }

int zlo_thumb_build(zl_ctx* ctx, int32_t id, int32_t spt) {
    Clip* c = ctx ? findClip(ctx, id) : nullptr; if (!c || spt < 1) return ZL_ERR_INVALID;
    buildThumb(c, spt); return ZL_OK;
}
int zlo_thumb_size(zl_ctx* ctx, int32_t id, int32_t* n, int32_t* nch) {
    Clip* c = ctx ? findClip(ctx, id) : nullptr; if (!c || !c->nThumb) return ZL_ERR_STATE;
    if (n) *n = c->nThumb;
    if (nch) *nch = c->nch;
    return ZL_OK;
}
int zlo_thumb_read_level0(zl_ctx* ctx, int32_t id, int8_t* out) {
    Clip* c = ctx ? findClip(ctx, id) : nullptr; if (!c || !c->nThumb || !out) return ZL_ERR_STATE;
    for (int ch = 0; ch < c->nch; ++ch) memcpy(out + (size_t) ch * (size_t) c->nThumb * 2, c->thumb[ch].data(), (size_t) c->nThumb * 2);
    return ZL_OK;
}
// AudioThumbnail::saveTo / loadFrom  juce_AudioThumbnail.cpp:593-639 (MinMaxValue::read/write :75-84: two raw bytes).
// juce_audio_utils does not build here, so this row is restated from the source lines only: parity unpinned.
int zlo_thumb_save_jatm(zl_ctx* ctx, int32_t id, void* out, uint64_t cap, uint64_t* size) {
    Clip* c = ctx ? findClip(ctx, id) : nullptr; if (!c || !size) return ZL_ERR_INVALID;
    if (!c->nThumb) buildThumb(c, 512);
    std::vector<unsigned char> o;
    auto writeInt = [&](int32_t v) { for (int i = 0; i < 4; ++i) o.push_back((unsigned char) ((uint32_t) v >> (8 * i))); };
    auto writeInt64 = [&](int64_t v) { for (int i = 0; i < 8; ++i) o.push_back((unsigned char) ((uint64_t) v >> (8 * i))); };
    o.insert(o.end(), {'j', 'a', 't', 'm'});
    writeInt(c->thumbSamples);
    writeInt64(c->length);                                                     // totalSamples
    writeInt64((int64_t) (c->length / c->thumbSamples) * c->thumbSamples);     // numSamplesFinished after the last setLevels (:735-743)
    writeInt(c->nThumb); writeInt(c->nch); writeInt((int) c->rate);
    writeInt64(0); writeInt64(0);
    for (int i = 0; i < c->nThumb; ++i)
        for (int ch = 0; ch < c->nch; ++ch) { o.push_back((unsigned char) c->thumb[ch][(size_t) i * 2]); o.push_back((unsigned char) c->thumb[ch][(size_t) i * 2 + 1]); }
    *size = o.size();
    if (!out || cap < o.size()) return ZL_ERR_STATE;
    memcpy(out, o.data(), o.size());
    return ZL_OK;
}
int zlo_thumb_load_jatm(zl_ctx* ctx, int32_t id, const void* jatm, uint64_t nbytes) {
    Clip* c = ctx ? findClip(ctx, id) : nullptr;
    const unsigned char* p = static_cast<const unsigned char*>(jatm);
    if (!c || !p || nbytes < 52 || p[0] != 'j' || p[1] != 'a' || p[2] != 't' || p[3] != 'm') return ZL_ERR_INVALID;
    auto readInt = [&](size_t at) { uint32_t v = 0; for (int i = 0; i < 4; ++i) v |= (uint32_t) p[at + (size_t) i] << (8 * i); return (int32_t) v; };
    auto readInt64 = [&](size_t at) { uint64_t v = 0; for (int i = 0; i < 8; ++i) v |= (uint64_t) p[at + (size_t) i] << (8 * i); return (int64_t) v; };
    const int spt = readInt(4), n = readInt(24), nch = readInt(28);
    if (spt < 1 || nch != c->nch || n != 1 + c->length / spt || readInt64(8) != c->length || nbytes < 52 + (uint64_t) n * (uint64_t) nch * 2) return ZL_ERR_INVALID;
    c->thumbSamples = spt; c->nThumb = n;
    for (int ch = 0; ch < nch; ++ch) c->thumb[ch].assign((size_t) n * 2, 0);
    size_t at = 52;
    for (int i = 0; i < n; ++i)
        for (int ch = 0; ch < nch; ++ch) { c->thumb[ch][(size_t) i * 2] = (int8_t) p[at++]; c->thumb[ch][(size_t) i * 2 + 1] = (int8_t) p[at++]; }
    return ZL_OK;
}
int zlo_thumb_pixels(zl_ctx* ctx, int32_t id, double t0, double t1, int32_t width, int8_t* out) {
    Clip* c = ctx ? findClip(ctx, id) : nullptr; if (!c || !out) return ZL_ERR_INVALID;
    if (!c->nThumb) buildThumb(c, 512);                                      // lib/WaveFormItem.cpp:22
    return refillCache(c, width, t0, t1, out) ? ZL_OK : ZL_ERR_INVALID;
}
int zlo_thumb_raster(zl_ctx* ctx, int32_t id, int32_t channel, double t0, double t1, int32_t width, int32_t height, float zoom, uint8_t* mask) {
    Clip* c = ctx ? findClip(ctx, id) : nullptr; if (!c || !mask || height < 1 || channel < 0 || channel >= c->nch) return ZL_ERR_INVALID;
    if (!c->nThumb) buildThumb(c, 512);
    std::vector<int8_t> px((size_t) c->nch * (size_t) std::max(width, 1) * 2);
    if (!refillCache(c, width, t0, t1, px.data())) return ZL_ERR_INVALID;
    memset(mask, 0, (size_t) width * (size_t) height);
    // CachedWindow::drawChannel  juce_AudioThumbnail.cpp:393-417, area = (0, 0, width, height)
    auto topY = (float) 0;
    auto bottomY = (float) height;
    auto midY = (topY + bottomY) * 0.5f;
    auto vscale = zoom * (bottomY - topY) / 256.0f;
    const int8_t* cacheData = px.data() + (size_t) channel * (size_t) width * 2;
    auto x = (float) 0;
    for (int w = width; --w >= 0;) {
        if (cacheData[1] > cacheData[0]) {                                   // isNonZero
            auto top = std::max(midY - cacheData[1] * vscale - 0.3f, topY);
            auto bottom = std::min(midY - cacheData[0] * vscale + 0.3f, bottomY);
            // QPainterContext::fillRectList -> QPainter::fillRect(int,int,int,int)  lib/QPainterContext.cpp:190-200
            const int rx = (int) x, ry = (int) top, rw = (int) 1.0f, rh = (int) (bottom - top);
            if (rw > 0 && rh > 0 && rx >= 0 && rx < width)
                for (int y = std::max(ry, 0); y < std::min(ry + rh, height); ++y) mask[(size_t) y * (size_t) width + (size_t) rx] = 1;
        }
        x += 1.0f;
        cacheData += 2;
    }
    return ZL_OK;
}

int zlo_passthrough_process(zl_ctx*, const zl_passthrough_params* p, const float* l, const float* r, int64_t n, float* out6) {
    if (!p || !l || !r || !out6 || n < 0) return ZL_ERR_INVALID;
    const size_t N = (size_t) n;
    passthroughProcess(*p, l, r, N, out6, out6 + N, out6 + 2 * N, out6 + 3 * N, out6 + 4 * N, out6 + 5 * N);
    return ZL_OK;
}
int zlo_levels_scan(zl_ctx*, const float* l, const float* r, int64_t n, int32_t* pa, int32_t* pb, double* sa, double* sb) {
    if (!l || !r || n < 0) return ZL_ERR_INVALID;
    if (pa) *pa = peakScan(l, (size_t) n);
    if (pb) *pb = peakScan(r, (size_t) n);
    if (sa) *sa = sumSquares(l, (size_t) n);
    if (sb) *sb = sumSquares(r, (size_t) n);
    return ZL_OK;
}
// the product's allocation-free variant (no counterpart in the reference): the same scan, refused beyond the fixed scratch
int zlo_levels_scan_static(zl_ctx* ctx, const float* l, const float* r, int64_t n, int32_t* pa, int32_t* pb, double* sa, double* sb) {
    if (n > ZL_STATIC_SCAN_MAX_FRAMES) return ZL_ERR_INVALID;
    return zlo_levels_scan(ctx, l, r, n, pa, pb, sa, sb);
}

// ---- rows n1 / n3: WAV in, WAV out — restated from the reference's reader and writer ----------------
// Reader: juce::WavAudioFormatReader ctor (juce_WavAudioFormat.cpp:933-1190) walks the chunks; readSamples/copySampleData
// (:1195-1245) left-justify into int32 (juce_AudioDataConverters.h:134,159,184,209); AudioFormatReader::read(AudioBuffer*)
// (juce_AudioFormatReader.cpp:139-183) scales with (float) i * (1.0f / 0x7fffffff); SamplerSynthSound keeps min(2, nch)
// channels (lib/SamplerSynthSound.cpp:40-46).  PINNED: tests/golden wav*_bytes / wav*_back come from the reference's own
// compiled reader and writer (oracle/ref/juce_ref.cpp ref_wav_roundtrip / ref_wav_read).
namespace {
struct WavIn {
    const unsigned char* p; uint64_t n; uint64_t at = 0;
    int byteAt(uint64_t i) const { return i < n ? p[i] : 0; }                  // InputStream reads past the end give zeros
    int64_t readLE(int bytes) { uint64_t v = 0; for (int i = 0; i < bytes; ++i) v |= (uint64_t) byteAt(at + (uint64_t) i) << (8 * i); at += (uint64_t) bytes; return (int64_t) v; }
    bool is(const char* fourcc, uint32_t v) const { return v == ((uint32_t) (unsigned char) fourcc[0] | ((uint32_t) (unsigned char) fourcc[1] << 8) | ((uint32_t) (unsigned char) fourcc[2] << 16) | ((uint32_t) (unsigned char) fourcc[3] << 24)); }
};
}  // namespace
int zlo_wav_probe(const void* wav, uint64_t nbytes, zl_wav_format* out) {
    if (!wav || !out) return ZL_ERR_INVALID;
    memset(out, 0, sizeof *out);
    WavIn in{static_cast<const unsigned char*>(wav), nbytes};
    uint64_t riffEnd = 0, dataLength = 0;
    bool isRF64 = false;
    const uint32_t first = (uint32_t) in.readLE(4);
    if (in.is("RF64", first)) { in.at += 4; isRF64 = true; }
    else if (in.is("RIFF", first)) { const uint64_t len = (uint32_t) in.readLE(4); riffEnd = len + in.at; }
    else return ZL_ERR_INVALID;
    const uint64_t startOfRIFFChunk = in.at;
    if (!in.is("WAVE", (uint32_t) in.readLE(4))) return ZL_ERR_INVALID;
    if (isRF64 && in.is("ds64", (uint32_t) in.readLE(4))) {
        const uint32_t length = (uint32_t) in.readLE(4);
        if (length < 28) return ZL_ERR_INVALID;
        const uint64_t chunkEnd = in.at + length + (length & 1);
        riffEnd = (uint64_t) in.readLE(8) + startOfRIFFChunk;
        dataLength = (uint64_t) in.readLE(8);
        in.at = chunkEnd;
    }
    unsigned numChannels = 0, bitsPerSample = 0;
    int bytesPerFrame = 0;
    double sampleRate = 0;
    bool usesFloat = false;
    int64_t lengthInSamples = 0;
    while (in.at < riffEnd && in.at < in.n) {
        const uint64_t chunkHeader = in.at;
        const uint32_t chunkType = (uint32_t) in.readLE(4);
        const uint32_t length = (uint32_t) in.readLE(4);
        uint64_t chunkEnd = in.at + length + (length & 1);
        if (in.is("fmt ", chunkType)) {
            const unsigned format = (unsigned short) in.readLE(2);
            numChannels = (unsigned) (int) (short) in.readLE(2);
            sampleRate = (double) (int) in.readLE(4);
            const int bytesPerSec = (int) in.readLE(4);
            in.at += 2;
            bitsPerSample = (unsigned) (int) (short) in.readLE(2);
            if (bitsPerSample > 64 && (int) sampleRate != 0) {
                bytesPerFrame = bytesPerSec / (int) sampleRate;
                if (numChannels != 0) bitsPerSample = 8 * (unsigned) bytesPerFrame / numChannels;
            } else bytesPerFrame = (int) (numChannels * bitsPerSample / 8);
            usesFloat = false;
            if (format == 3) usesFloat = true;
            else if (format == 0xfffe) {
                if (length < 40) bytesPerFrame = 0;
                else {
                    in.at += 4;                                                  // size and bitsPerSample
                    in.at += 4;                                                  // channel mask
                    unsigned char guid[16];
                    for (int i = 0; i < 16; ++i) guid[i] = (unsigned char) in.readLE(1);
                    static const unsigned char pcm[16] = {1, 0, 0, 0, 0, 0, 0x10, 0, 0x80, 0, 0, 0xaa, 0, 0x38, 0x9b, 0x71};
                    static const unsigned char flt[16] = {3, 0, 0, 0, 0, 0, 0x10, 0, 0x80, 0, 0, 0xaa, 0, 0x38, 0x9b, 0x71};
                    static const unsigned char amb[16] = {1, 0, 0, 0, 0x21, 0x07, 0xd3, 0x11, 0x86, 0x44, 0xc8, 0xc1, 0xca, 0, 0, 0};
                    if (!memcmp(guid, flt, 16)) usesFloat = true;
                    else if (memcmp(guid, pcm, 16) && memcmp(guid, amb, 16)) bytesPerFrame = 0;
                }
            } else if (format == 0x674f || format == 0x6750 || format == 0x6751 || format == 0x676f || format == 0x6770 || format == 0x6771) return ZL_ERR_INVALID;
            else if (format != 1) bytesPerFrame = 0;
        } else if (in.is("data", chunkType)) {
            if (isRF64) { if (dataLength > 0) chunkEnd = in.at + dataLength + (dataLength & 1); }
            else dataLength = length;
            out->data_offset = in.at; out->data_bytes = dataLength;
            lengthInSamples = bytesPerFrame > 0 ? (int64_t) (dataLength / (uint64_t) bytesPerFrame) : 0;
        } else {
            static const char* const known[] = {"bext", "smpl", "inst", "INST", "cue ", "axml", "LIST", "acid", "Trkn"};
            bool isKnown = false;
            for (const char* k : known) isKnown = isKnown || in.is(k, chunkType);
            if (!isKnown && chunkEnd <= in.at) break;
        }
        // A ds64 data length near 2^64 wraps chunkEnd back into the file (the reference's int64 sum does the same and its walk
        // never ends); the boundary ends the walk there: a chunk must end after its own header.
        if (chunkEnd <= chunkHeader) break;
        in.at = chunkEnd;
    }
    out->channels = (int32_t) numChannels; out->bits_per_sample = (int32_t) bitsPerSample; out->is_float = usesFloat ? 1 : 0;
    out->bytes_per_frame = bytesPerFrame; out->sample_rate = sampleRate; out->frames = lengthInSamples;
    // WavAudioFormat::createReaderFor :1718-1731, then the formats copySampleData knows :1232-1245
    if (!(sampleRate > 0) || (int) numChannels <= 0 || bytesPerFrame <= 0 || bitsPerSample > 32) return ZL_ERR_INVALID;
    if (!(bitsPerSample == 8 || bitsPerSample == 16 || bitsPerSample == 24 || bitsPerSample == 32) || (usesFloat && bitsPerSample != 32)) return ZL_ERR_INVALID;
    if ((int) (numChannels * (bitsPerSample / 8)) > bytesPerFrame) return ZL_ERR_INVALID;
    return ZL_OK;
}
int zlo_clip_create_from_wav(zl_ctx* ctx, const void* wav, uint64_t nbytes, int32_t* clip_id) {
    zl_wav_format f;
    if (!ctx || !wav || !clip_id || zlo_wav_probe(wav, nbytes, &f) != ZL_OK) return ZL_ERR_INVALID;
    if (f.frames < 1 || f.frames > 2147483647LL - 8) return ZL_ERR_INVALID;
    const int keep = std::min(2, (int) f.channels), bps = f.bits_per_sample / 8;
    std::vector<std::vector<float>> pcm((size_t) keep, std::vector<float>((size_t) f.frames));
    WavIn in{static_cast<const unsigned char*>(wav), nbytes};
    const float scale = 1.0f / static_cast<float>(0x7fffffff);                  // juce_AudioFormatReader.cpp:41
    for (int64_t fr = 0; fr < f.frames; ++fr)
        for (int ch = 0; ch < keep; ++ch) {
            const uint64_t q = f.data_offset + (uint64_t) fr * (uint64_t) f.bytes_per_frame + (uint64_t) (ch * bps);
            uint32_t raw = 0;
            for (int i = 0; i < bps; ++i) raw |= (uint32_t) in.byteAt(q + (uint64_t) i) << (8 * i);
            float v;
            if (f.is_float) memcpy(&v, &raw, 4);
            else {
                int32_t fixed;
                if (bps == 1) fixed = (int32_t) ((raw - 128u) << 24);           // UInt8::getAsInt32LE
                else if (bps == 2) fixed = (int32_t) (raw << 16);               // Int16
                else if (bps == 3) fixed = (int32_t) (raw << 8);                // Int24
                else fixed = (int32_t) raw;
                v = (float) fixed * scale;
            }
            pcm[(size_t) ch][(size_t) fr] = v;
        }
    const float* planar[2] = {pcm[0].data(), keep > 1 ? pcm[1].data() : nullptr};
    return zlo_clip_create_from_pcm(ctx, planar, keep, f.frames, f.sample_rate, clip_id);
}
int zlo_clip_create_from_wav_file(zl_ctx* ctx, const char* path, int32_t* clip_id) {
    if (!ctx || !path || !clip_id) return ZL_ERR_INVALID;
    FILE* fp = fopen(path, "rb");
    if (!fp) return ZL_ERR_INVALID;
    std::vector<unsigned char> bytes;
    unsigned char buf[65536];
    size_t got;
    while ((got = fread(buf, 1, sizeof buf, fp)) > 0) bytes.insert(bytes.end(), buf, buf + got);
    fclose(fp);
    return zlo_clip_create_from_wav(ctx, bytes.data(), bytes.size(), clip_id);
}
int zlo_clip_read_pcm(zl_ctx* ctx, int32_t id, float* const* planar, int32_t nch, int64_t length) {
    Clip* c = ctx ? findClip(ctx, id) : nullptr;
    if (!c || !planar || nch != c->nch || length != c->length) return ZL_ERR_INVALID;
    for (int ch = 0; ch < nch; ++ch) memcpy(planar[ch], c->data[ch].data(), (size_t) length * 4);
    return ZL_OK;
}
// Writer: AudioFormatWriter::writeFromFloatArrays -> convertFloatsToInts (juce_AudioFormatWriter.cpp:64-79,160-185), then
// WavAudioFormatWriter::write keeps the top bits (juce_WavAudioFormat.cpp:1336-1369) behind writeHeader's layout (:1391-1500).
static int wavEncode(const float* l, const float* r, size_t N, double rate, int bits, void* out, uint64_t cap, uint64_t* size) {
    const int nch = r ? 2 : 1, bytesPerSample = bits / 8, bytesPerFrame = nch * bytesPerSample;
    const uint64_t audioDataSize = (uint64_t) bytesPerFrame * N;
    const uint64_t total = 104 + audioDataSize + (audioDataSize & 1);
    *size = total;
    if (!out || cap < total) return ZL_ERR_STATE;
    unsigned char* o = static_cast<unsigned char*>(out);
    size_t at = 0;
    auto put = [&](uint64_t v, int bytes) { for (int i = 0; i < bytes; ++i) o[at++] = (unsigned char) (v >> (8 * i)); };
    auto fourcc = [&](const char* t) { memcpy(o + at, t, 4); at += 4; };
    int64_t riffChunkSize = (int64_t) (4 + 8 + 40 + 8 + audioDataSize + (audioDataSize & 1) + (8 + 28));
    riffChunkSize += (riffChunkSize & 1);
    fourcc("RIFF"); put((uint64_t) riffChunkSize, 4); fourcc("WAVE");
    fourcc("JUNK"); put(28 + 24, 4); for (int i = 0; i < 28 + 24; ++i) o[at++] = 0;
    fourcc("fmt "); put(16, 4); put(1, 2); put((uint64_t) nch, 2); put((uint64_t) (int) rate, 4);
    put((uint64_t) (int) ((double) bytesPerFrame * rate), 4); put((uint64_t) bytesPerFrame, 2); put((uint64_t) bits, 2);
    fourcc("data"); put(audioDataSize, 4);
    for (size_t j = 0; j < N; ++j)
        for (int ch = 0; ch < nch; ++ch) {
            const double samp = (ch ? r : l)[j];
            int fixed;
            if (samp <= -1.0) fixed = std::numeric_limits<int>::min();
            else if (samp >= 1.0) fixed = std::numeric_limits<int>::max();
            else fixed = roundToInt(std::numeric_limits<int>::max() * samp);
            put((uint64_t) (uint32_t) (fixed >> (32 - bits)), bytesPerSample);
        }
    if (audioDataSize & 1) o[at++] = 0;
    return ZL_OK;
}
int zlo_wav_encode(zl_ctx* ctx, const float* l, const float* r, int64_t n, double rate, int32_t bits, void* out, uint64_t cap, uint64_t* size) {
    if (!ctx || !l || n < 1 || !(rate > 0) || (bits != 16 && bits != 24) || !size) return ZL_ERR_INVALID;
    return wavEncode(l, r, (size_t) n, rate, bits, out, cap, size);
}
int zlo_bounce_wav(zl_ctx* ctx, int32_t s, int32_t b, int32_t which, int32_t bits, void* out, uint64_t cap, uint64_t* size) {
    if (!ctx || !size || (bits != 16 && bits != 24) || s < 0 || s >= ctx->cfg.n_sessions || b >= ctx->cfg.buses_per_session || (b >= 0 && (which < 0 || which > 3)))
        return ZL_ERR_INVALID;
    const size_t N = (size_t) ctx->nblocks * (size_t) ctx->nframes;
    const float* l = nullptr;
    if (b < 0) { if ((ctx->cfg.outputs & ZL_OUT_MASTER) && !ctx->master.empty()) l = ctx->master.data() + (size_t) s * 2 * N; }
    else if ((ctx->cfg.outputs & (ZL_OUT_RAW << which)) && !ctx->bus[which].empty()) l = ctx->bus[which].data() + ((size_t) s * (size_t) ctx->cfg.buses_per_session + (size_t) b) * 2 * N;
    if (!l || N < 1) return ZL_ERR_STATE;
    return wavEncode(l, l + N, N, ctx->cfg.sample_rate, bits, out, cap, size);
}

// ---- row n2: SyncTimer's step ring, restated from lib/SyncTimer.cpp ----------------------------------------
// StepData :42-77, SyncTimerPrivate (ring set-up :279-290, delayedStep :346-362, hiResTimerCallback :392-415, process :452-702,
// updateScheduleAheadAmount :712-715), SyncTimer::start/stop :870-924, setBpm :955-975, queueClipToStart/StopOnChannel :816-860,
// scheduleClipCommand :1011-1048.  Pinned against the reference's own compiled SyncTimer.cpp (oracle/ref/synctimer_ref.cpp).
namespace {
struct OStepData {
    std::vector<zl_clip_command> clipCommands;
    std::vector<std::array<int, 3>> timerCommands;            // (operation, parameter, parameter2)
    OStepData* previous = nullptr; OStepData* next = nullptr;
    uint64_t index = 0;
    bool played = true;
    void ensureFresh() { if (played) { played = false; timerCommands.clear(); clipCommands.clear(); } }
};
const uint64_t OStepRingCount = 32768, OBeatSubdivisions = 96, ONanosecondsPerMinute = 60000000000ull;
uint64_t oSubbeatCountToNanoseconds(const uint64_t& bpm, const uint64_t& subBeatCount) { return (subBeatCount * ONanosecondsPerMinute) / (bpm * OBeatSubdivisions); }
float oNanosecondsToSubbeatCount(const uint64_t& bpm, const uint64_t& nanoseconds) { return nanoseconds / (ONanosecondsPerMinute / (bpm * OBeatSubdivisions)); }
}  // namespace
struct zl_sched {
    std::vector<OStepData> stepRing = std::vector<OStepData>(OStepRingCount);
    OStepData* stepReadHead = nullptr;
    uint64_t stepNextPlaybackPosition = 0, jackPlayhead = 0, jackNextPlaybackPosition = 0, jackSubbeatLengthInMicroseconds = 0, jackLatency = 0;
    uint64_t stepReadHeadOnStart = 0, cumulativeBeat = 0, scheduleAheadAmount = 0, timerBpm = 120;
    double jackPlayheadBpm = 120;
    int beat = 0;
    bool isPaused = true;
    zl_sched() {
        OStepData* previous = &stepRing[OStepRingCount - 1];
        for (uint64_t i = 0; i < OStepRingCount; ++i) { stepRing[i].index = i; previous->next = &stepRing[i]; stepRing[i].previous = previous; previous = &stepRing[i]; }
        stepReadHead = &stepRing[0];
    }
    OStepData* delayedStep(uint64_t delay) {
        uint64_t step = 0;
        if (isPaused) step = (stepReadHead->index + delay + 1) % OStepRingCount;
        else step = (stepReadHeadOnStart + std::max(cumulativeBeat + delay, jackPlayhead + 1)) % OStepRingCount;
        OStepData* stepData = &stepRing[step];
        stepData->ensureFresh();
        return stepData;
    }
    void updateScheduleAheadAmount() { scheduleAheadAmount = (oNanosecondsToSubbeatCount(timerBpm, jackLatency * (float) 1000000)) + 1; }
    void setBpm(uint64_t bpm) {
        if (timerBpm != bpm) {
            timerBpm = bpm;
            jackSubbeatLengthInMicroseconds = oSubbeatCountToNanoseconds(timerBpm, 1) / 1000;
            updateScheduleAheadAmount();
            delayedStep(0)->timerCommands.push_back({10 /* SetBpmOperation */, (int) bpm, 0});
        }
    }
    void scheduleClipCommand(const zl_clip_command& command, uint64_t delay) {
        OStepData* stepData = delayedStep(delay);
        bool foundExisting = false;
        for (zl_clip_command& existingCommand : stepData->clipCommands) {
            if (equivalentTo(existingCommand, command)) {
                if (command.change_looping) { existingCommand.looping = command.looping; existingCommand.change_looping = true; }
                if (command.change_pitch) { existingCommand.pitch_change = command.pitch_change; existingCommand.change_pitch = true; }
                if (command.change_speed) { existingCommand.speed_ratio = command.speed_ratio; existingCommand.change_speed = true; }
                if (command.change_gain_db) { existingCommand.gain_db = command.gain_db; existingCommand.change_gain_db = true; }
                if (command.change_volume) { existingCommand.volume = command.volume; existingCommand.change_volume = true; }
                if (command.start_playback) existingCommand.start_playback = true;
                foundExisting = true;
            }
        }
        if (!foundExisting) stepData->clipCommands.push_back(command);
    }
};
int zlo_sched_create(double sample_rate, int32_t buffer_size, zl_sched** out) {
    if (!out || !(sample_rate > 0) || buffer_size < 1) return ZL_ERR_INVALID;
    zl_sched* d = new zl_sched;
    d->jackSubbeatLengthInMicroseconds = oSubbeatCountToNanoseconds(d->timerBpm, 1) / 1000;
    d->jackLatency = (1000 * (double) buffer_size) / (double) sample_rate;
    d->updateScheduleAheadAmount();
    *out = d; return ZL_OK;
}
int zlo_sched_destroy(zl_sched* d) { if (!d) return ZL_ERR_INVALID; delete d; return ZL_OK; }
int zlo_sched_start(zl_sched* d, int32_t bpm) {
    if (!d || bpm < 1) return ZL_ERR_INVALID;
    d->setBpm((uint64_t) bpm);
    d->stepReadHeadOnStart = d->stepReadHead->index;
    d->isPaused = false;
    return ZL_OK;
}
int zlo_sched_stop(zl_sched* d) {
    if (!d) return ZL_ERR_INVALID;
    d->isPaused = true;
    d->beat = 0; d->cumulativeBeat = 0; d->jackPlayhead = 0;
    // :893-917, literally: walk the whole ring from the read head; every command still waiting has its volume forced to 0 and is
    // re-scheduled at delay 0 - isPaused is already true here (pausedChanged is a direct connection, :750-752), so that is the
    // step after the read head.  The read head's own commands move there; that step then meets its own commands (merged into
    // themselves) and is marked played; the first later command finds it played, clears it, and the rest pile up there, muted,
    // to come out when the read head next passes.
    for (uint64_t step = 0; step < OStepRingCount; ++step) {
        OStepData* stepData = &d->stepRing[(step + d->stepReadHead->index) % OStepRingCount];
        if (!stepData->played) {
            for (size_t i = 0; i < stepData->clipCommands.size(); ++i) {
                zl_clip_command c = stepData->clipCommands[i];
                c.change_volume = 1; c.volume = 0.0f;
                stepData->clipCommands[i] = c;
                d->scheduleClipCommand(c, 0);
            }
            stepData->played = true;
        }
    }
    return ZL_OK;
}
int zlo_sched_set_bpm(zl_sched* d, int32_t bpm) { if (!d || bpm < 1) return ZL_ERR_INVALID; d->setBpm((uint64_t) bpm); return ZL_OK; }
int zlo_sched_schedule(zl_sched* d, const zl_clip_command* cmd, uint64_t delay) { if (!d || !cmd) return ZL_ERR_INVALID; d->scheduleClipCommand(*cmd, delay); return ZL_OK; }
int zlo_sched_schedule_channel_enabled(zl_sched* d, uint64_t delay, int32_t channel, int32_t enabled) {
    if (!d) return ZL_ERR_INVALID;
    d->delayedStep(delay)->timerCommands.push_back({8 /* SamplerChannelEnabledStateOperation */, channel, enabled});
    return ZL_OK;
}
int zlo_sched_queue_start(zl_sched* d, int32_t clip, int32_t midiChannel) {
    if (!d) return ZL_ERR_INVALID;
    zl_clip_command command; zlo_clip_command_default(&command);
    command.clip = clip; command.midi_channel = midiChannel; command.midi_note = 60; command.change_volume = true; command.volume = 1.0;
    command.looping = true; command.stop_playback = true; command.start_playback = true;
    const uint64_t nextZeroBeat = d->isPaused ? 0 : (OBeatSubdivisions * 4) - (d->cumulativeBeat % (OBeatSubdivisions * 4));
    d->scheduleClipCommand(command, d->cumulativeBeat + nextZeroBeat < d->jackPlayhead ? nextZeroBeat + OBeatSubdivisions * 4 : nextZeroBeat);
    return ZL_OK;
}
int zlo_sched_queue_stop(zl_sched* d, int32_t clip, int32_t midiChannel) {
    if (!d) return ZL_ERR_INVALID;
    for (uint64_t step = 0; step < OStepRingCount; ++step) {
        OStepData* stepData = &d->stepRing[step];
        if (!stepData->played)
            for (size_t i = 0; i < stepData->clipCommands.size(); ++i)
                if (stepData->clipCommands[i].clip == clip) { stepData->clipCommands.erase(stepData->clipCommands.begin() + (long) i); break; }
    }
    zl_clip_command command; zlo_clip_command_default(&command);
    command.clip = clip; command.midi_channel = midiChannel; command.midi_note = 60; command.stop_playback = true;
    d->delayedStep(0)->clipCommands.push_back(command);
    return ZL_OK;
}
int zlo_sched_advance(zl_sched* d, int32_t nblocks, int32_t nframesI, zl_block_clock* clocks, zl_sched_event* events, int32_t cap, int32_t* count) {
    if (!d || nblocks < 1 || nframesI < 1 || !clocks || !count || cap < 0 || (cap && !events)) return ZL_ERR_INVALID;
    std::vector<zl_sched_event> out;
    const uint32_t nframes = (uint32_t) nframesI;
    for (int b = 0; b < nblocks; ++b) {
        const uint64_t current_usecs = clocks[b].current_usecs, next_usecs = clocks[b].next_usecs;
        const uint64_t microsecondsPerFrame = (next_usecs - current_usecs) / nframes;
        double thisStepBpm{d->jackPlayheadBpm};
        double thisStepSubbeatLengthInMicroseconds{double(oSubbeatCountToNanoseconds(d->jackPlayheadBpm, 1)) / 1000.0};
        if (!d->isPaused) {
            if (d->jackPlayhead == 0) d->jackNextPlaybackPosition = current_usecs;
        }
        if (d->stepNextPlaybackPosition == 0) d->stepNextPlaybackPosition = current_usecs;
        uint32_t firstAvailableFrame{0}, relativePosition{0};
        while (d->stepNextPlaybackPosition < next_usecs && firstAvailableFrame < nframes) {
            OStepData* stepData = d->stepReadHead;
            d->stepReadHead = d->stepReadHead->next;
            if (d->stepNextPlaybackPosition <= current_usecs) { relativePosition = firstAvailableFrame; ++firstAvailableFrame; }
            else {
                relativePosition = std::clamp<uint32_t>((d->stepNextPlaybackPosition - current_usecs) / microsecondsPerFrame, firstAvailableFrame, nframes - 1);
                firstAvailableFrame = relativePosition;
            }
            if (!stepData->played) {
                for (const zl_clip_command& clipCommand : stepData->clipCommands) {
                    zl_sched_event ev; memset(&ev, 0, sizeof ev);
                    ev.kind = ZL_SCHED_CLIP_COMMAND; ev.block = b; ev.tick = d->jackPlayhead; ev.cmd = clipCommand;
                    out.push_back(ev);
                }
                for (const auto& command : stepData->timerCommands) {
                    if (command[0] == 8) {
                        zl_sched_event ev; memset(&ev, 0, sizeof ev);
                        ev.kind = ZL_SCHED_CHANNEL_ENABLED; ev.block = b; ev.tick = d->jackPlayhead; ev.channel = command[1]; ev.enabled = command[2];
                        out.push_back(ev);
                    } else if (command[0] == 10) {
                        const uint64_t newBpm{std::clamp<uint64_t>(uint64_t(command[1]), 50, 200)};
                        d->setBpm(newBpm);
                        thisStepBpm = newBpm;
                    }
                }
                stepData->played = true;
            }
            if (d->jackPlayheadBpm != thisStepBpm) {
                d->jackPlayheadBpm = thisStepBpm;
                thisStepSubbeatLengthInMicroseconds = oSubbeatCountToNanoseconds(d->jackPlayheadBpm, 1) / 1000;
            }
            if (!d->isPaused) {
                ++d->jackPlayhead;
                d->jackNextPlaybackPosition += thisStepSubbeatLengthInMicroseconds;
            }
            d->stepNextPlaybackPosition += thisStepSubbeatLengthInMicroseconds;
        }
        clocks[b].playhead = d->isPaused ? d->stepReadHead->index : d->jackPlayhead;                      // :990-1009
        clocks[b].playhead_usecs = d->isPaused ? d->stepNextPlaybackPosition : d->jackNextPlaybackPosition;
        clocks[b].subbeat_usecs = d->jackSubbeatLengthInMicroseconds;
        if (!d->isPaused)                                                                                  // hiResTimerCallback
            while (d->cumulativeBeat < (d->jackPlayhead + (d->scheduleAheadAmount * 2))) { d->beat = (d->beat + 1) % (OBeatSubdivisions * 4); ++d->cumulativeBeat; }
    }
    *count = (int32_t) out.size();
    if ((int32_t) out.size() > cap) return ZL_ERR_STATE;
    if (!out.empty()) memcpy(events, out.data(), out.size() * sizeof(zl_sched_event));
    return ZL_OK;
}
int zlo_sched_state(zl_sched* d, uint64_t* playhead, uint64_t* playhead_usecs, uint64_t* subbeat_usecs, uint64_t* cumulative_beat, int32_t* beat, uint64_t* schedule_ahead) {
    if (!d) return ZL_ERR_INVALID;
    if (playhead) *playhead = d->isPaused ? d->stepReadHead->index : d->jackPlayhead;
    if (playhead_usecs) *playhead_usecs = d->isPaused ? d->stepNextPlaybackPosition : d->jackNextPlaybackPosition;
    if (subbeat_usecs) *subbeat_usecs = d->jackSubbeatLengthInMicroseconds;
    if (cumulative_beat) *cumulative_beat = d->cumulativeBeat;
    if (beat) *beat = d->beat;
    if (schedule_ahead) *schedule_ahead = d->scheduleAheadAmount;
    return ZL_OK;
}

// One mix spanning GPUs: the oracle is one process; a group of one reduces to itself, larger groups are summed by the test
// harness (tests/test_shard_gloo.py) in rank order.
int zlo_comm_unique_id(void* id128) { if (!id128) return ZL_ERR_INVALID; memset(id128, 0, 128); return ZL_OK; }
int zlo_comm_init(zl_ctx* ctx, const void* id, int32_t rank, int32_t nranks) { return (ctx && id && nranks >= 1 && rank >= 0 && rank < nranks) ? ZL_OK : ZL_ERR_INVALID; }
int zlo_mix_reduce(zl_ctx* ctx, int32_t) { return ctx ? ZL_OK : ZL_ERR_INVALID; }
int zlo_comm_destroy(zl_ctx* ctx) { return ctx ? ZL_OK : ZL_ERR_INVALID; }
int zlo_host_alloc(void** p, size_t bytes) { if (!p) return ZL_ERR_INVALID; *p = malloc(bytes ? bytes : 1); return *p ? ZL_OK : ZL_ERR_NOMEM; }
int zlo_host_free(void* p) { free(p); return ZL_OK; }

// ---- oracle-only probes used to pin the restatement against the real JUCE objects --------------
// ADSR run: noteOn, n_on samples, noteOff, remaining samples. renote != 0: noteOff before every
// sample of the release part (libzl Q3, lib/SamplerSynthVoice.cpp:253-256).
ZL_EXPORT int zlo_adsr_run(double sr, float a, float d, float s, float r, int n_on, int n_total, int renote, float* out) {
    Adsr e; e.setSampleRate(sr); e.setParameters(a, d, s, r); e.noteOn();
    for (int i = 0; i < n_total; ++i) {
        if (i == n_on || (renote && i > n_on)) e.noteOff();
        out[i] = e.getNextSample();
    }
    return ZL_OK;
}
ZL_EXPORT int zlo_round_to_int(double v) { return roundToInt(v); }
ZL_EXPORT void zlo_set_float(float mn, float mx, int8_t* out) { setFloat(mn, mx, out); }

} // extern "C"
