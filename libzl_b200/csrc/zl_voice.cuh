// zl_voice.cuh — per-voice state machine (exact, frame-by-frame recurrences, register-resident between tiles)
// and the in-kernel SamplerChannel::handleCommand.  Used by the render kernel in zl_render.cu.
#pragma once
#include "zl_device.cuh"

namespace zl {

// SamplerChannel::handleCommand   lib/SamplerSynth.cpp:187-230  (voices = the bus's V voice records)
__device__ inline void bus_handle_command(VoiceDev* voices, int V, int busMidiChannel, const CmdDev& cd,
                                          const RenderParams& P) {
    const zl_clip_command& c = cd.c;
    const int sound = (c.clip >= 0 && c.clip < P.nClips && P.clips[c.clip].valid) ? c.clip : -1;
    if (busMidiChannel != c.midi_channel) return;
    if (c.stop_playback || c.start_playback) {
        if (c.stop_playback) {
            for (int i = 0; i < V; ++i) {
                VoiceDev& v = voices[i];
                if (v.sound == sound && (v.flags & VF_HAS_CMD) && cmd_equivalent(v.cmd, c)) voice_stop(v, true);
            }
        }
        if (c.start_playback) {
            for (int i = 0; i < V; ++i) {
                VoiceDev& v = voices[i];
                if (!(v.flags & VF_PLAYING)) {
                    voice_set_command(v, c, P.clips, P.sliceBounds);
                    v.startTick = cd.tick;
                    // juce::Synthesiser::startVoice  JUCE/juce_audio_basics/synthesisers/juce_Synthesiser.cpp:306-328
                    if (sound >= 0) {
                        if (v.sound >= 0) voice_stop(v, false);
                        v.sound = sound;
                        v.midiNote = c.midi_note;
                        voice_start(v, sound, P);
                    }
                    break;
                }
            }
        }
    } else {
        for (int i = 0; i < V; ++i) {
            VoiceDev& v = voices[i];
            if (v.sound == sound && (v.flags & VF_HAS_CMD) && cmd_equivalent(v.cmd, c))
                voice_set_command(v, c, P.clips, P.sliceBounds);
        }
    }
}


// ---------------------------------------------------------------------------------------------
// Register-resident part of a voice record: what the frame loop of SamplerSynthVoice::process
// (lib/SamplerSynthVoice.cpp:197-262) reads or writes.  The lane that owns the voice keeps it for the
// whole launch; the cold fields (the ClipCommand copy, start tick, note) stay in the HBM record and are
// touched only by bus_handle_command.
struct VoiceHot {
    double p, ratio, relDen, srcLen;
    uint64_t nextLoopTick, nextLoopUsecs;
    float env, rRate, aRate, dRate, sustain, release;
    float gain;               // lgain == rgain (both are always set together: :131-132, :77-78)
    float posGain, posProgress;
    int32_t adsrState, flags, sound, clip, slice, looping;
};

__device__ __forceinline__ void hot_load(VoiceHot& h, const VoiceDev& v) {
    h.p = v.p; h.ratio = v.ratio; h.relDen = v.relDen; h.srcLen = v.srcLen;
    h.nextLoopTick = v.nextLoopTick; h.nextLoopUsecs = v.nextLoopUsecs;
    h.env = v.env; h.rRate = v.rRate; h.aRate = v.aRate; h.dRate = v.dRate; h.sustain = v.sustain; h.release = v.release;
    h.gain = v.lgain; h.posGain = v.posGain; h.posProgress = v.posProgress;
    h.adsrState = v.adsrState; h.flags = v.flags; h.sound = v.sound; h.clip = v.clip;
    h.slice = v.cmd.slice; h.looping = v.cmd.looping;
}

__device__ __forceinline__ void hot_store(const VoiceHot& h, VoiceDev& v) {
    v.p = h.p; v.nextLoopTick = h.nextLoopTick; v.nextLoopUsecs = h.nextLoopUsecs;
    v.env = h.env; v.rRate = h.rRate;
    v.posGain = h.posGain; v.posProgress = h.posProgress;
    v.adsrState = h.adsrState; v.flags = h.flags; v.sound = h.sound; v.clip = h.clip;
}

// What the gather needs to know about one voice for one block (shared memory, 32 bytes = 2 x LDS.128).
struct __align__(16) VoiceGather {
    const float* data;        // interleaved source frames
    int32_t sampleDuration;   // length - 1                              lib/SamplerSynthVoice.cpp:191
    int32_t nch;
    float gain, vol, lPan, rPan;
};

// Per-block constants of the owner lane's walk (registers).
struct VoiceBlockCtl {
    double stopD, startD, relStart, usPerFrame;
    uint64_t lengthInTicks;
    int32_t fWrap;            // beat-aligned loop: first frame >= the last wrap whose clock passes nextLoopUsecs
    int32_t beatAligned;
    int32_t spanStart, lastPos;   // algorithmic source-span accounting
    uint32_t span;
    int32_t nch;
    int32_t releaseArmed;     // Q3 fixed: noteOff() once per block instead of on every frame of the release window
};

// First frame f in [from, nframes) with  current_usecs + (u64)(f * usPerFrame) >= nextLoopUsecs
// (lib/SamplerSynthVoice.cpp:232), nframes if none.  The left side is non-decreasing in f, so the per-frame
// test of the reference collapses to one comparison  frame == fWrap.
__device__ __forceinline__ bool loop_clock_hit(const zl_block_clock& ck, double usPerFrame, uint64_t nextLoopUsecs, int f) {
    return ck.current_usecs + d2u64(__dmul_rn((double) (uint32_t) f, usPerFrame)) >= nextLoopUsecs;
}
__device__ inline int first_wrap_frame(const zl_block_clock& ck, double usPerFrame, uint64_t nextLoopUsecs, int from, int nframes) {
    if (from >= nframes) return nframes;
    if (loop_clock_hit(ck, usPerFrame, nextLoopUsecs, from)) return from;
    if (!loop_clock_hit(ck, usPerFrame, nextLoopUsecs, nframes - 1)) return nframes;
    int lo = from, hi = nframes - 1;          // invariant: !hit(lo), hit(hi)
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (loop_clock_hit(ck, usPerFrame, nextLoopUsecs, mid)) hi = mid; else lo = mid;
    }
    return hi;
}

// Start of one block for the owner lane (the per-block prelude of process(), :176-196 / J: juce_Sampler.cpp:119-126).
// Returns true when the voice renders in this block; fills the walk constants and the gather record.
template <bool JMODE>
__device__ __forceinline__ bool voice_block_begin(VoiceHot& h, const RenderParams& P, const zl_block_clock& ck, int nframes,
                                                  VoiceBlockCtl& b, VoiceGather& g) {
    if (h.sound < 0) return false;
    const ClipDev& snd = P.clips[h.sound];
    g.data = snd.data; g.nch = snd.nch; g.sampleDuration = snd.length - 1; g.gain = h.gain;
    b.nch = snd.nch; b.releaseArmed = 0; b.span = 0; b.spanStart = (int) h.p; b.lastPos = b.spanStart;
    if (JMODE) {
        g.vol = 1.0f; g.lPan = 0.0f; g.rPan = 0.0f;
        b.stopD = (double) snd.length;                                        // juce_Sampler.cpp:155
        return true;
    }
    if (!(snd.length > 0 && (h.flags & VF_HAS_CMD))) return false;            // :176-177
    if (h.clip < 0) return false;
    const ClipDev& c = P.clips[h.clip];
    if (h.nextLoopUsecs == 0) {                                               // :179-182
        const uint64_t diff = h.nextLoopTick - ck.playhead;
        h.nextLoopUsecs = ck.playhead_usecs + diff * ck.subbeat_usecs;
    }
    b.usPerFrame = (double) ((ck.next_usecs - ck.current_usecs) / (uint64_t) nframes);   // :183
    b.stopD = (double) clip_stop_pos(snd, P.sliceBounds, h.slice);
    b.startD = (double) clip_start_pos(c, P.sliceBounds, h.slice);
    b.relStart = __dsub_rn(b.stopD, __dmul_rn((double) h.release, snd.rate));             // :253
    b.beatAligned = c.beatAligned; b.lengthInTicks = c.lengthInTicks;
    b.fWrap = (h.looping && c.beatAligned) ? first_wrap_frame(ck, b.usPerFrame, h.nextLoopUsecs, 0, nframes) : nframes;
    g.vol = c.volume; g.lPan = c.lPan; g.rPan = c.rPan;
    return true;
}

// `n` frames of the block starting at block frame `f0`, frame by frame, exactly as the reference steps them:
// writes (pos, alpha, env) of each rendered frame at tp/ta/te[i] and returns how many were rendered;
// `active` goes false when the voice ends inside the tile.   lib/SamplerSynthVoice.cpp:197-262
template <bool Q3>
__device__ __forceinline__ int voice_walk_tile_z(VoiceHot& h, VoiceBlockCtl& b, const zl_block_clock& ck, int f0, int n, int nframes,
                                                 int* tp, float* ta, float* te, bool& active) {
    int i = 0;
    for (; i < n; ) {
        const int pos = (int) h.p;
        tp[i] = pos;
        ta[i] = __double2float_rn(__dsub_rn(h.p, (double) pos));
        te[i] = adsr_next(h);
        const int frame = f0 + i;
        ++i;
        b.lastPos = pos;
        h.p = __dadd_rn(h.p, h.ratio);                                        // :223
        if (h.looping) {
            bool wrap = false;
            if (b.beatAligned) {
                if (frame == b.fWrap) {                                       // :232-236
                    h.nextLoopTick = h.nextLoopTick + b.lengthInTicks;
                    const uint64_t diff = h.nextLoopTick - ck.playhead;
                    h.nextLoopUsecs = ck.playhead_usecs + diff * ck.subbeat_usecs;
                    b.fWrap = first_wrap_frame(ck, b.usPerFrame, h.nextLoopUsecs, frame + 1, nframes);
                    wrap = true;
                }
            } else if (h.p >= b.stopD) {                                      // :243-247
                wrap = true;
            }
            if (wrap) {
                h.p = b.startD;
                if ((int) b.startD < pos) {                                   // backward jump: close the span, open a new one
                    b.span += (uint32_t) (pos + 2 - b.spanStart); b.spanStart = (int) b.startD; b.lastPos = b.spanStart - 2;
                }
            }
        } else {
            if (h.p >= b.stopD) { voice_stop(h, false); active = false; return i; }    // :249-251
            else if (h.p >= b.relStart) {                                     // :253-256 (Q3: re-armed on every frame)
                if (Q3 || !b.releaseArmed) { adsr_note_off(h); b.releaseArmed = 1; }
            }
        }
        if (h.adsrState == kIdle) { voice_stop(h, false); active = false; return i; }  // :258-261
    }
    return i;
}

// J-mode: juce::SamplerVoice::renderNextBlock   juce_Sampler.cpp:128-161
__device__ __forceinline__ int voice_walk_tile_j(VoiceHot& h, VoiceBlockCtl& b, int n, int* tp, float* ta, float* te, bool& active) {
    int i = 0;
    for (; i < n; ) {
        const int pos = (int) h.p;
        tp[i] = pos;
        ta[i] = __double2float_rn(__dsub_rn(h.p, (double) pos));
        te[i] = adsr_next(h);
        ++i;
        b.lastPos = pos;
        h.p = __dadd_rn(h.p, h.ratio);
        if (h.p > b.stopD) { voice_stop(h, false); active = false; return i; }
    }
    return i;
}

}  // namespace zl
