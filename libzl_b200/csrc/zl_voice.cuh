// zl_voice.cuh — per-voice state machine steps (exact, frame-by-frame form) and the in-kernel
// SamplerChannel::handleCommand.  Used by the render kernels in zl_render.cu.
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

// Per-block constants the gather needs for one voice.
struct VoiceBlock {
    const float* data;
    int32_t nch;
    int32_t sampleDuration;   // length - 1                              lib/SamplerSynthVoice.cpp:191
    float lgain, rgain, vol, lPan, rPan;
    int32_t nrend;            // frames rendered in this block (loop ran for frames [0, nrend))
    uint32_t srcFloats;       // source floats spanned in this block (unique samples incl. the +1 tap)
};

// One block of SamplerSynthVoice::process (lib/SamplerSynthVoice.cpp:174-270), frame by frame, writing
// the per-frame (pos, alpha, env) the gather consumes.  Returns frames rendered.
__device__ inline int voice_walk_z(VoiceDev& v, const RenderParams& P, const zl_block_clock& ck, int nframes,
                                   int* tpos, float* talpha, float* tenv, VoiceBlock& vb) {
    vb.nrend = 0; vb.srcFloats = 0;
    if (v.sound < 0) return 0;
    const ClipDev& snd = P.clips[v.sound];
    if (!(snd.length > 0 && (v.flags & VF_HAS_CMD))) return 0;
    if (v.clip < 0) return 0;
    const ClipDev& c = P.clips[v.clip];
    if (v.nextLoopUsecs == 0) {                                               // :179-182
        const uint64_t diff = v.nextLoopTick - ck.playhead;
        v.nextLoopUsecs = ck.playhead_usecs + diff * ck.subbeat_usecs;
    }
    const double usPerFrame = (double) ((ck.next_usecs - ck.current_usecs) / (uint64_t) nframes);   // :183
    const int stopPosition = clip_stop_pos(snd, P.sliceBounds, v.cmd.slice);
    const int startPosition = clip_start_pos(c, P.sliceBounds, v.cmd.slice);
    const double stopD = (double) stopPosition;
    const double relStart = __dsub_rn(stopD, __dmul_rn((double) v.release, snd.rate));   // :253
    const bool looping = v.cmd.looping != 0;
    const bool q3 = (P.quirks & ZL_Q3_RENOTEOFF) != 0;
    vb.data = snd.data; vb.nch = snd.nch; vb.sampleDuration = snd.length - 1;
    vb.lgain = v.lgain; vb.rgain = v.rgain; vb.vol = c.volume; vb.lPan = c.lPan; vb.rPan = c.rPan;
    bool releaseArmed = false;
    int rendered = 0;
    int spanStart = (int) v.p, lastPos = spanStart;
    uint32_t span = 0;
    for (int frame = 0; frame < nframes; ++frame) {
        const int pos = (int) v.p;
        if (pos < lastPos) { span += (uint32_t) (lastPos + 2 - spanStart); spanStart = pos; }   // loop wrap
        lastPos = pos;
        tpos[frame] = pos;
        talpha[frame] = __double2float_rn(__dsub_rn(v.p, (double) pos));
        tenv[frame] = adsr_next(v);
        ++rendered;
        v.p = __dadd_rn(v.p, v.ratio);                                        // :223
        if (looping) {
            if (c.beatAligned) {
                if (ck.current_usecs + d2u64(__dmul_rn((double) (uint32_t) frame, usPerFrame)) >= v.nextLoopUsecs) {   // :232
                    v.nextLoopTick = v.nextLoopTick + c.lengthInTicks;
                    const uint64_t diff = v.nextLoopTick - ck.playhead;
                    v.nextLoopUsecs = ck.playhead_usecs + diff * ck.subbeat_usecs;
                    v.p = (double) startPosition;
                }
            } else if (v.p >= stopD) {
                v.p = (double) startPosition;
            }
        } else {
            if (v.p >= stopD) { voice_stop(v, false); break; }
            else if (v.p >= relStart) {
                if (q3 || !releaseArmed) { adsr_note_off(v); releaseArmed = true; }
            }
        }
        if (v.adsrState == kIdle) { voice_stop(v, false); break; }
    }
    span += (uint32_t) (lastPos + 2 - spanStart);
    vb.srcFloats = span * (uint32_t) snd.nch;
    vb.nrend = rendered;
    return rendered;
}

// One block of juce::SamplerVoice::renderNextBlock (juce_Sampler.cpp:117-163).
__device__ inline int voice_walk_j(VoiceDev& v, const RenderParams& P, int nframes,
                                   int* tpos, float* talpha, float* tenv, VoiceBlock& vb) {
    vb.nrend = 0; vb.srcFloats = 0;
    if (v.sound < 0) return 0;
    const ClipDev& snd = P.clips[v.sound];
    vb.data = snd.data; vb.nch = snd.nch; vb.sampleDuration = snd.length - 1;
    vb.lgain = v.lgain; vb.rgain = v.rgain; vb.vol = 1.0f; vb.lPan = 0.0f; vb.rPan = 0.0f;
    const double lenD = (double) snd.length;
    int rendered = 0;
    const int spanStart = (int) v.p;
    int lastPos = spanStart;
    for (int frame = 0; frame < nframes; ++frame) {
        const int pos = (int) v.p;
        lastPos = pos;
        tpos[frame] = pos;
        talpha[frame] = __double2float_rn(__dsub_rn(v.p, (double) pos));
        tenv[frame] = adsr_next(v);
        ++rendered;
        v.p = __dadd_rn(v.p, v.ratio);
        if (v.p > lenD) { voice_stop(v, false); break; }
    }
    vb.srcFloats = (uint32_t) (lastPos + 2 - spanStart) * (uint32_t) snd.nch;
    vb.nrend = rendered;
    return rendered;
}

}  // namespace zl
