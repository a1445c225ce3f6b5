// zl_device.cuh — device-resident tables of the sampler voice-render path (sm_100a).
//
// The voice table replaces SamplerSynth's 12 x 8 SamplerSynthVoice objects (lib/SamplerSynth.cpp:23,258;
// private state lib/SamplerSynthVoice.cpp:20-39), the clip table replaces ClipAudioSource +
// SamplerSynthSound (lib/ClipAudioSource.cpp:63-82, lib/SamplerSynthSound.cpp:28-59), the bus table
// replaces SamplerChannel + JackPassthroughPrivate (lib/SamplerSynth.cpp:33-61, lib/JackPassthrough.cpp:27-31).
// All audio arithmetic uses explicit round-to-nearest intrinsics (__fmul_rn/__fadd_rn/...) so nvcc never
// contracts a*b+c into an FMA: results are the ones a -ffp-contract=off build of the reference gives.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/zl_b200.h"

namespace zl {

enum AdsrState : int { kIdle = 0, kAttack = 1, kDecay = 2, kSustain = 3, kRelease = 4 };

// Voice flags
enum : int {
    VF_PLAYING = 1,      // SamplerSynthVoice::isPlaying            lib/SamplerSynthVoice.h:32
    VF_HAS_CMD = 2,      // d->clipCommand != nullptr
    VF_HAS_POSID = 4,    // d->clipPositionId > -1
};

struct ClipDev {
    const float* data;       // interleaved frames (float2 for stereo), length + 8 frames, zero padded
    double rate;             // sourceSampleRate
    double srcLen;           // getDuration() * rate                  lib/SamplerSynthVoice.cpp:120
    double relDen;           // release * rate (f32 * f64): ADSR noteOff denominator   juce_ADSR.h:138
    int32_t nch;
    int32_t length;
    float lengthInBeats;
    float volume;            // volumeAbsolute
    float lPan, rPan;        // 0.5*(1+pan), 0.5*(1-pan) in f64 -> f32 lib/SamplerSynthVoice.cpp:193-194
    float sustain, release;
    float aRate, dRate, rRate;   // juce::ADSR::recalculateRates at sampleRate = rate   juce_ADSR.h:230-247
    int32_t root;
    int32_t beatAligned;     // trunc(lengthInBeats) == lengthInBeats  lib/SamplerSynthVoice.cpp:227
    int32_t sliceOff;        // into the slice-bounds table: [sliceOff + 2*(slice+1)] = start, +1 = stop
    int32_t nSliceEntries;   // entries for slice = -1 .. nSliceEntries-2
    uint64_t lengthInTicks;  // (quint64)(lengthInBeats * 96)          lib/SamplerSynthVoice.cpp:234
    int32_t valid;
    int32_t pad;
};

struct VoiceDev {
    double p;                // sourceSamplePosition
    double ratio;            // pitchRatio
    double srcLen;           // sourceSampleLength
    uint64_t startTick, nextLoopTick, nextLoopUsecs;
    double relDen;           // ADSR release * sampleRate (f32 * f64) JUCE/juce_audio_basics/utilities/juce_ADSR.h:138
    float env, rRate;        // ADSR envelopeVal, releaseRate (recomputed by noteOff)
    float aRate, dRate, sustain, release;   // ADSR parameters/rates captured by setParameters at startNote
    float lgain, rgain;
    float posGain, posProgress;
    int32_t adsrState;
    int32_t flags;
    int32_t sound;           // getCurrentlyPlayingSound() clip id or -1
    int32_t clip;            // d->clip id or -1
    int32_t midiNote;
    int32_t pad;
    zl_clip_command cmd;     // *d->clipCommand (68 bytes)
    int32_t pad2;
};
static_assert(sizeof(VoiceDev) % 8 == 0, "VoiceDev must be 8-byte granular");

struct BusDev {
    zl_passthrough_params pt;
    int32_t enabled;
    int32_t cmdBegin, cmdCount;   // this launch's commands of the bus, sorted by block (stable)
};

struct CmdDev {
    zl_clip_command c;
    int32_t block;
    uint64_t tick;
};

struct TileDesc;
struct VoiceGather;

struct RenderParams {
    VoiceDev* voices;        // [nbus][V]
    const ClipDev* clips;
    const int32_t* sliceBounds;
    BusDev* buses;           // [nbus]
    const CmdDev* cmds;
    const int32_t* cmdRange; // [nbus][2] = (first command, count) of this launch per bus, or nullptr: no commands
    const zl_block_clock* clocks;   // [nblocks]
    const double* pow2;      // pow(2, k/12), k = -127..127 (host libm, like the reference's std::pow)
    float* out[4];           // raw, dry, fx1, fx2 planes [nbus][2][stride] or nullptr
    float* master;           // [nsess][2][stride] or nullptr
    int32_t* peaks;          // [nblocks][nbus][2] or nullptr
    double* sumsq;           // [nblocks][nbus][2] or nullptr
    unsigned long long* voiceFrames;   // global counter
    double outRate;
    int64_t stride;          // floats per plane row (>= nblocks*nframes)
    int32_t nSessions, B, V, nClips;
    int32_t nblocks, nframes;
    int32_t mode, quirks;
    // per-block hand-over between plan_kernel and gather_kernel (scratch, overwritten every block)
    TileDesc* descs;         // [tilesPerBlock][nVoices]
    VoiceGather* vgs;        // [nVoices]
    float* te;               // [tilesPerBlock][nVoices][kT] envelope rows (only frames >= TileDesc.ct are used)
    double* ps;              // [tilesPerBlock][nVoices][kT] position rows (only kSlotBit voice-tiles)
    int32_t* walkQueue;      // [nVoices] voices deferred by plan_kernel to walk_kernel in this block
    int32_t* walkCount;      // [nblocks][2] queue lengths: serial walkers (front of the queue), others (back); zeroed per launch
    int32_t nVoices, tilesPerBlock, blk, planBusesPerCta;
    int32_t planGroups, planCtaCap, walkCtaCap;     // bus groups of plan_kernel; grid caps beside a running gather (0 = none)
    int32_t dbgFlags;                               // profiling aid (ZL_B200_DEBUG_SKIP bit 4): plan_kernel queues nothing, walk_kernel runs empty
};

// ---------------------------------------------------------------------------------------------
// float / double -> uint64 as the reference's deployment target does it (aarch64 fcvtzu saturates,
// negative -> 0).  PTX cvt.rzi.u64 has the same semantics.   lib/SamplerSynthVoice.cpp:123,234
__device__ __forceinline__ uint64_t f2u64(float x) { return __float2ull_rz(x); }
__device__ __forceinline__ uint64_t d2u64(double x) { return __double2ull_rz(x); }

// juce::roundToInt: low word of (double)v + 2^52 + 2^51   JUCE/juce_core/maths/juce_MathsFunctions.h:471-485
__device__ __forceinline__ int juce_round_to_int(double v) {
    return __double2loint(__dadd_rn(v, 6755399441055744.0));
}

// ---------------------------------------------------------------------------------------------
// juce::ADSR on the voice record   JUCE/juce_audio_basics/utilities/juce_ADSR.h
template <class VT> __device__ __forceinline__ void adsr_reset(VT& v) { v.env = 0.0f; v.adsrState = kIdle; }          // :107-111

template <class VT> __device__ __forceinline__ void adsr_set_parameters(VT& v, const ClipDev& c) {                    // :76-83 (state is idle here)
    v.aRate = c.aRate; v.dRate = c.dRate; v.rRate = c.rRate; v.sustain = c.sustain; v.release = c.release;
    v.relDen = c.relDen;
}

template <class VT> __device__ __forceinline__ void adsr_note_on(VT& v) {                           // :114-129
    if (v.aRate > 0.0f) v.adsrState = kAttack;
    else if (v.dRate > 0.0f) { v.env = 1.0f; v.adsrState = kDecay; }
    else v.adsrState = kSustain;
}

template <class VT> __device__ __forceinline__ void adsr_note_off(VT& v) {                          // :132-146
    if (v.adsrState != kIdle) {
        if (v.release > 0.0f) {
            v.rRate = __double2float_rn(__ddiv_rn((double) v.env, v.relDen));
            v.adsrState = kRelease;
        } else {
            adsr_reset(v);
        }
    }
}

template <class VT> __device__ __forceinline__ void adsr_next_state(VT& v) {                        // :249-257
    if (v.adsrState == kAttack) v.adsrState = (v.dRate > 0.0f ? kDecay : kSustain);
    else if (v.adsrState == kDecay) v.adsrState = kSustain;
    else if (v.adsrState == kRelease) adsr_reset(v);
}

template <class VT> __device__ __forceinline__ float adsr_next(VT& v) {                             // :153-191
    const int st = v.adsrState;
    if (st == kIdle) return 0.0f;
    if (st == kAttack) {
        v.env = __fadd_rn(v.env, v.aRate);
        if (v.env >= 1.0f) { v.env = 1.0f; adsr_next_state(v); }
    } else if (st == kDecay) {
        v.env = __fsub_rn(v.env, v.dRate);
        if (v.env <= v.sustain) { v.env = v.sustain; adsr_next_state(v); }
    } else if (st == kSustain) {
        v.env = v.sustain;
    } else {
        v.env = __fsub_rn(v.env, v.rRate);
        if (v.env <= 0.0f) adsr_next_state(v);
    }
    return v.env;
}

// ---------------------------------------------------------------------------------------------
// slice bounds: (int)(getStartPosition(slice)*rate), (int)(getStopPosition(slice)*rate), tabulated on the
// host when the clip parameters are set (lib/ClipAudioSource.cpp:261-277, lib/SamplerSynthSound.cpp:96-104).
__device__ __forceinline__ int clip_start_pos(const ClipDev& c, const int32_t* sb, int slice) {
    const int e = (slice > -1 && slice + 1 < c.nSliceEntries) ? slice + 1 : 0;
    return sb[c.sliceOff + 2 * e];
}
__device__ __forceinline__ int clip_stop_pos(const ClipDev& c, const int32_t* sb, int slice) {
    const int e = (slice > -1 && slice + 1 < c.nSliceEntries) ? slice + 1 : 0;
    return sb[c.sliceOff + 2 * e + 1];
}

// ClipCommand::equivalentTo   lib/ClipCommand.h:33-39
__device__ __forceinline__ bool cmd_equivalent(const zl_clip_command& a, const zl_clip_command& b) {
    return a.clip == b.clip
        && ((a.change_slice && b.change_slice && a.slice == b.slice)
            || (!a.change_slice && !b.change_slice && a.midi_note == b.midi_note && a.midi_channel == b.midi_channel));
}

// SamplerSynthVoice::stopNote   lib/SamplerSynthVoice.cpp:146-169  (J: juce_Sampler.cpp:100-111)
template <class VT> __device__ __forceinline__ void voice_stop(VT& v, bool allowTailOff) {
    if (allowTailOff) {
        adsr_note_off(v);
    } else {
        v.sound = -1;
        adsr_reset(v);
        v.clip = -1;
        v.flags = 0;
        v.nextLoopTick = 0;
        v.nextLoopUsecs = 0;
    }
}

// SamplerSynthVoice::setCurrentCommand   lib/SamplerSynthVoice.cpp:58-98
__device__ __forceinline__ void voice_set_command(VoiceDev& v, const zl_clip_command& c, const ClipDev* clips,
                                                  const int32_t* sb) {
    if (v.flags & VF_HAS_CMD) {
        if (c.change_looping) { v.cmd.looping = c.looping; v.cmd.change_looping = 1; }
        if (c.change_pitch) { v.cmd.pitch_change = c.pitch_change; v.cmd.change_pitch = 1; }
        if (c.change_speed) { v.cmd.speed_ratio = c.speed_ratio; v.cmd.change_speed = 1; }
        if (c.change_gain_db) { v.cmd.gain_db = c.gain_db; v.cmd.change_gain_db = 1; }
        if (c.change_volume) {
            v.cmd.volume = c.volume; v.cmd.change_volume = 1;
            v.lgain = c.volume; v.rgain = c.volume;
        }
        if (c.change_slice) v.cmd.slice = c.slice;
        if (c.start_playback && v.sound >= 0 && v.clip >= 0)
            v.p = (double) clip_start_pos(clips[v.clip], sb, v.cmd.slice);
    } else {
        v.cmd = c;
        v.flags |= VF_HAS_CMD;
    }
    if (v.flags & VF_HAS_CMD) v.flags |= VF_PLAYING; else v.flags &= ~VF_PLAYING;
}

// startNote: Z lib/SamplerSynthVoice.cpp:110-144, J juce_Sampler.cpp:78-98
__device__ __forceinline__ void voice_start(VoiceDev& v, int clipId, const RenderParams& P) {
    const ClipDev& c = P.clips[clipId];
    const int k = min(127, max(-127, v.cmd.midi_note - c.root));
    const double ratio = __ddiv_rn(__dmul_rn(P.pow2[k + 127], c.rate), P.outRate);
    const float velocity = v.cmd.volume;
    if (P.mode == ZL_MODE_J) {
        v.ratio = ratio;
        v.p = 0.0;
        v.lgain = velocity; v.rgain = velocity;
        adsr_set_parameters(v, c);
        adsr_note_on(v);
        v.clip = clipId;
        v.srcLen = c.srcLen;
    } else if (c.length > 0) {
        v.ratio = ratio;
        v.clip = clipId;
        v.srcLen = c.srcLen;
        v.p = (double) clip_start_pos(c, P.sliceBounds, v.cmd.slice);
        v.nextLoopTick = f2u64(__fadd_rn(__ull2float_rn(v.startTick), __fmul_rn(c.lengthInBeats, 96.0f)));
        v.nextLoopUsecs = 0;
        v.flags |= VF_HAS_POSID;
        v.lgain = velocity; v.rgain = velocity;
        adsr_reset(v);
        adsr_set_parameters(v, c);
        adsr_note_on(v);
    }
}

}  // namespace zl
