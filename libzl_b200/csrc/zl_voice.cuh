// zl_voice.cuh — per-voice state machine (exact, frame-by-frame recurrences, register-resident between tiles)
// and the in-kernel SamplerChannel::handleCommand.  Used by the render kernel in zl_render.cu.
#pragma once
#include <limits.h>
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
    v.posProgress = h.posProgress;      // posGain is written where it is produced (plan_voice_block / gather_kernel), never copied back
    v.adsrState = h.adsrState; v.flags = h.flags; v.sound = h.sound; v.clip = h.clip;
}

// What the gather needs to know about one voice for one block (shared memory, 32 bytes = 2 x LDS.128).
struct __align__(16) VoiceGather {
    const float* data;        // interleaved source frames
    int32_t sampleDuration;   // length - 1                              lib/SamplerSynthVoice.cpp:191
    int32_t nch;
    float gain, vol, lPan, rPan;
};

// What the gather needs to know about one voice for one tile (64 bytes = 4 x LDG.128), written by plan_kernel.
struct __align__(16) TileDesc {
    double p0a, Ra, p0b, Rb;         // position segments a: frames [0, cb), b: frames [cb, nrend)
    float e0a, Ea, e0b, Eb;          // envelope segments a: frames [0, ce), b: frames [ce, ct); frames >= ct: table row
    int32_t cb, ce, ct;
    int32_t nrend;                   // frames of the tile this voice renders | flag bits (zl_render.cu)
};
static_assert(sizeof(TileDesc) == 64, "TileDesc is 64 bytes");
static_assert(sizeof(VoiceGather) == 32, "VoiceGather is 32 bytes");

// Per-block constants of the owner lane's walk (registers).
struct VoiceBlockCtl {
    double stopD, startD, relStart, usPerFrame;
    uint64_t lengthInTicks;
    int32_t fWrap;            // beat-aligned loop: first frame >= the last wrap whose clock passes nextLoopUsecs
    int32_t beatAligned;
    int32_t spanStart, lastPos;   // algorithmic source-span accounting
    uint32_t span;
    int32_t nch, clipFrames;
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
    if (!snd.valid) return false;                                             // destroyed clip: zl_clip_destroy stops its voices; belt and braces
    g.data = snd.data; g.nch = snd.nch; g.sampleDuration = snd.length - 1; g.gain = h.gain;
    b.nch = snd.nch; b.clipFrames = snd.length; b.releaseArmed = 0; b.span = 0; b.spanStart = (int) h.p; b.lastPos = b.spanStart;
    if (JMODE) {
        g.vol = 1.0f; g.lPan = 0.0f; g.rPan = 0.0f;
        // juce_Sampler.cpp:155 stops when p > length; stored as the next double up so every mode tests p >= stopD
        b.stopD = __longlong_as_double(__double_as_longlong((double) snd.length) + 1);
        b.startD = 0.0; b.relStart = 0.0; b.beatAligned = 0; b.fWrap = nframes;
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


// ---------------------------------------------------------------------------------------------
// Closed-form position segments.
//
// The reference advances the source position with  p = RN(p + ratio)  in f64 once per frame (:223).  While p
// stays inside one binade [2^e, 2^(e+1)) every p is a multiple of u = 2^(e-52), so RN(p + ratio) = p + R with
// R = ratio rounded to a multiple of u — the SAME R for every step (ties excepted: |ratio - R| = u/2 rounds by
// the parity of p).  Hence  p_k = p_0 + k*R  exactly, and a DFMA evaluates it exactly because the true value is
// representable.  A tile of frames is therefore described by at most two (p0, R) segments, broken at binade
// crossings and loop wraps, where ONE real DADD + the reference's event tests are executed; the rare tiles
// that need more (p < 1, ties, tiny loops) fall back to the per-frame walk below.  Bit-exact by construction.
struct PosPlan {
    double p0a, Ra, p0b, Rb;
    int32_t cb;               // first tile frame of segment b (>= n: unused)
    int32_t nPos;             // frames the position logic renders in this tile (voice stops after frame nPos-1 if stopped)
    int32_t kRel;             // Z one-shot: first increment k (after tile frame k-1) with p_k >= relStart; INT_MAX: none
    int32_t stopped;          // the position logic ended the voice (:249-251 / juce_Sampler.cpp:155-159)
    int32_t irregular;        // not expressible: use the per-frame walk
};

// R for the binade of p, or false when the closed form does not apply.
__device__ __forceinline__ bool binade_step(double p, double ratio, double& R, double& top) {
    if (!(p >= 1.0)) return false;
    const int hi = __double2hiint(p);
    top = __hiloint2double((hi & 0x7ff00000) + 0x00100000, 0);            // 2^(e+1)
    if (!(ratio < __dmul_rn(top, 0.25))) return false;                       // binade shorter than ~2 frames
    const double M = __dmul_rn(top, 0.75);                                  // 1.5 * 2^e: ulp(M) = u
    R = __dsub_rn(__dadd_rn(M, ratio), M);
    const double u = __dmul_rn(top, 1.1102230246251565e-16);                // 2^(e+1) * 2^-53 = u
    if (__dmul_rn(fabs(__dsub_rn(ratio, R)), 2.0) == u) return false;       // tie: rounding would depend on p's parity
    return R > 0.0;
}

// Smallest k in [1, kmax] with p + k*R >= lim, or kmax + 1.  p + k*R is exact while < top (see above).
__device__ __forceinline__ int first_k_reaching(double p, double R, double lim, int kmax) {
    if (__fma_rn((double) kmax, R, p) < lim) return kmax + 1;
    int k = (int) __ddiv_rn(__dsub_rn(lim, p), R);
    k = max(1, min(kmax, k));
    while (k > 1 && __fma_rn((double) (k - 1), R, p) >= lim) --k;
    while (k < kmax && __fma_rn((double) k, R, p) < lim) ++k;
    return k;
}

// The increment after block frame `frame` whose position was `plast`: the reference's :223-251 (J: :154-159) for
// one frame, position side only.  Returns true when the voice ends here.
template <bool JMODE>
__device__ __forceinline__ bool pos_explicit_step(VoiceHot& h, VoiceBlockCtl& b, const zl_block_clock& ck, double plast, int frame,
                                                  int nframes, int kInc, int& kRel) {
    const int pos = (int) plast;
    b.lastPos = pos;
    double pn = __dadd_rn(plast, h.ratio);
    if (!JMODE && h.looping) {
        bool wrap = false;
        if (b.beatAligned) {
            if (frame == b.fWrap) {                                           // :232-236
                h.nextLoopTick = h.nextLoopTick + b.lengthInTicks;
                const uint64_t diff = h.nextLoopTick - ck.playhead;
                h.nextLoopUsecs = ck.playhead_usecs + diff * ck.subbeat_usecs;
                b.fWrap = first_wrap_frame(ck, b.usPerFrame, h.nextLoopUsecs, frame + 1, nframes);
                wrap = true;
            }
        } else if (pn >= b.stopD) wrap = true;                                // :243-247
        if (wrap) {
            pn = b.startD;
            if ((int) b.startD < pos) {                                       // backward jump: close the span, open a new one
                b.span += (uint32_t) (pos + 2 - b.spanStart); b.spanStart = (int) b.startD; b.lastPos = b.spanStart - 2;
            }
        }
        h.p = pn;
        return false;
    }
    h.p = pn;
    if (pn >= b.stopD) return true;                                           // :249-251 / J :155-159
    if (!JMODE && kRel == INT_MAX && pn >= b.relStart) kRel = kInc;           // :253
    return false;
}

// Plans the positions of tile frames [0, n) (block frames f0 ..) from h.p; advances h.p / loop clock / span.
template <bool JMODE>
__device__ __forceinline__ void plan_positions(VoiceHot& h, VoiceBlockCtl& b, const zl_block_clock& ck, int f0, int n, int nframes,
                                               PosPlan& pl) {
    pl.cb = n; pl.nPos = n; pl.kRel = INT_MAX; pl.stopped = 0; pl.irregular = 0;
    pl.p0a = h.p; pl.Ra = 0.0; pl.p0b = 0.0; pl.Rb = 0.0;
    const bool looping = !JMODE && h.looping;
    const bool oneShotZ = !JMODE && !h.looping;
    int c = 0, nseg = 0;
    while (c < n) {
        double R, top;
        if (nseg == 2 || !binade_step(h.p, h.ratio, R, top)) { pl.irregular = 1; return; }
        const double p = h.p;
        if (nseg == 0) { pl.p0a = p; pl.Ra = R; } else { pl.p0b = p; pl.Rb = R; pl.cb = c; }
        ++nseg;
        const int mMax = n - c;
        int mCap = mMax;
        if (looping && b.beatAligned) mCap = min(mCap, b.fWrap - (f0 + c) + 1);      // the wrap frame ends its segment
        const bool wrapsOrStops = !(looping && b.beatAligned);                          // p_k >= stopD is an event
        const double lim = wrapsOrStops ? fmin(top, b.stopD) : top;
        const double q = __fma_rn((double) mCap, R, p);
        if (q < lim && mCap == mMax) {
            // the whole rest of the tile is one event-free run; q is exactly the next tile's first position
            if (oneShotZ && pl.kRel == INT_MAX && q >= b.relStart) pl.kRel = c + first_k_reaching(p, R, b.relStart, mMax);
            b.lastPos = (int) __fma_rn((double) (mMax - 1), R, p);
            h.p = q;
            return;
        }
        const int k1 = first_k_reaching(p, R, lim, mCap);
        const int m = min(k1, mCap);                                                    // frames in this segment; >= 1
        if (oneShotZ && pl.kRel == INT_MAX && m > 1 && __fma_rn((double) (m - 1), R, p) >= b.relStart)
            pl.kRel = c + first_k_reaching(p, R, b.relStart, m - 1);
        const double plast = __fma_rn((double) (m - 1), R, p);
        if (pos_explicit_step<JMODE>(h, b, ck, plast, f0 + c + m - 1, nframes, c + m, pl.kRel)) {
            pl.nPos = c + m; pl.stopped = 1;
            return;
        }
        c += m;
    }
}

// Fallback: the same tile, frame by frame, writing each frame's position to tpd[i] (shared-memory slot).
template <bool JMODE>
__device__ __forceinline__ void walk_positions(VoiceHot& h, VoiceBlockCtl& b, const zl_block_clock& ck, int f0, int n, int nframes,
                                               double* tpd, PosPlan& pl) {
    pl.cb = n; pl.nPos = n; pl.kRel = INT_MAX; pl.stopped = 0; pl.irregular = 1;
    for (int i = 0; i < n; ++i) {
        const double p = h.p;
        tpd[i] = p;
        if (pos_explicit_step<JMODE>(h, b, ck, p, f0 + i, nframes, i + 1, pl.kRel)) { pl.nPos = i + 1; pl.stopped = 1; return; }
    }
}

// ---------------------------------------------------------------------------------------------
// Closed-form envelope segments: the same binade argument as for the position, in f32.  While juce::ADSR attacks,
// decays or (single noteOff) releases, env +- rate is exact-linear as long as env stays in one f32 binade, so a tile
// is described by up to two (e0, E) segments  env(c) = e0 + (c - cs) * E  (E = 0: sustain / J-mode idle), evaluated
// exactly by one FFMA; what does not fit (the first frames of an attack from 0, state changes beyond two segments,
// the per-frame re-armed release of quirk Q3) is stepped frame by frame into the shared-memory row from frame ct on.
struct EnvPlan {
    float e0a, Ea, e0b, Eb;
    int32_t ce;               // first tile frame of segment b
    int32_t ct;               // first tile frame read from the table row (INT_MAX: none)
};

// E = rate rounded to the ulp of e0's binade, or false when the closed form does not apply (tiny e0, ties, rate too big).
__device__ __forceinline__ bool binade_step_f(float e0, float rate, float& E, float& top, float& bottom) {
    if (!(e0 >= 1.0e-30f) || !(rate > 0.0f)) return false;
    const int bits = __float_as_int(e0);
    bottom = __int_as_float(bits & 0x7f800000);
    top = __fmul_rn(bottom, 2.0f);
    if (!(rate < __fmul_rn(top, 0.25f))) return false;
    const float M = __fmul_rn(top, 0.75f);                                  // 1.5 * 2^e: ulp(M) = u
    E = __fsub_rn(__fadd_rn(M, rate), M);
    const float u = __fmul_rn(top, 5.9604644775390625e-08f);               // 2^(e+1) * 2^-24
    if (__fmul_rn(fabsf(__fsub_rn(rate, E)), 2.0f) == u) return false;      // tie: rounding would depend on env's parity
    return E > 0.0f;
}

// Largest x in [0, X] with e0 + x*E < lim (up) or e0 - x*E >= lim (down); x = 0 always qualifies by construction.
__device__ __forceinline__ int env_run_up(float e0, float E, float lim, int X) {
    if (__fmaf_rn((float) X, E, e0) < lim) return X;
    int x = (int) __fdiv_rn(__fsub_rn(lim, e0), E);
    x = max(0, min(X, x));
    while (x > 0 && !(__fmaf_rn((float) x, E, e0) < lim)) --x;
    while (x < X && __fmaf_rn((float) (x + 1), E, e0) < lim) ++x;
    return x;
}
__device__ __forceinline__ int env_run_down(float e0, float E, float lim, int X) {
    if (__fmaf_rn((float) X, -E, e0) >= lim) return X;
    int x = (int) __fdiv_rn(__fsub_rn(e0, lim), E);
    x = max(0, min(X, x));
    while (x > 0 && !(__fmaf_rn((float) x, -E, e0) >= lim)) --x;
    while (x < X && __fmaf_rn((float) (x + 1), -E, e0) >= lim) ++x;
    return x;
}

// juce::ADSR::noteOff with the f64 division replaced by a multiply with 1/relDen, kept bit-exact: the product is
// within 2 ulp(f64) of the quotient, so the f32 result can differ only when it sits next to an f32 rounding
// midpoint; those (2^-26 of the cases) take the real division.   juce_ADSR.h:132-146
__device__ __forceinline__ void adsr_note_off_fast(VoiceHot& v, double relRcp) {
    if (v.adsrState == kIdle) return;
    if (!(v.release > 0.0f)) { adsr_reset(v); return; }
    const double q = __dmul_rn((double) v.env, relRcp);
    const unsigned lo = (unsigned) __double2loint(q) & 0x1fffffffu;
    float rr;
    if (lo - 0x0ffffff8u <= 16u || !(q > 1.0e-30)) rr = __double2float_rn(__ddiv_rn((double) v.env, v.relDen));
    else rr = __double2float_rn(q);
    v.rRate = rr;
    v.adsrState = kRelease;
}

// Envelope of tile frames [0, nPos), with the release logic of :253-261, in two parts so the warp can hand out the few
// shared-memory table rows only to the voices that need one.
//
// Part 1 - closed-form segments.  Returns the frames rendered (<= nPos) when the tile's envelope is complete
// (resume = -1; `ended` is set when the ADSR going idle ends the voice, :258-261).  Otherwise resume = the first tile
// frame that has to be stepped frame by frame (plan_envelope_walk) and the return value is meaningless.
template <bool JMODE, bool Q3>
__device__ __forceinline__ int plan_envelope_closed(VoiceHot& h, VoiceBlockCtl& b, const PosPlan& pl, EnvPlan& ep, bool& ended, int& resume) {
    ended = false; resume = -1;
    const int nPos = pl.nPos;
    const int lastNoteOffInc = pl.stopped ? nPos - 1 : nPos;      // the increment that stops the voice fires no noteOff
    ep.e0a = 0.0f; ep.Ea = 0.0f; ep.e0b = 0.0f; ep.Eb = 0.0f; ep.ce = nPos; ep.ct = INT_MAX;
    int i = 0, nseg = 0;
    while (i < nPos) {
        // next increment at which noteOff fires (after frame k0 - 1)
        int k0 = INT_MAX;
        if (!JMODE && pl.kRel <= lastNoteOffInc && (Q3 || !b.releaseArmed)) { k0 = max(pl.kRel, i + 1); if (k0 > lastNoteOffInc) k0 = INT_MAX; }
        const int runEnd = min(nPos, k0);
        const int st = h.adsrState;
        if (nseg == 2 || (!JMODE && st == kIdle) || (st == kAttack && h.env == 0.0f)) break;   // the rest goes frame by frame
        if (!JMODE && Q3 && st == kRelease && k0 == i + 1) break;  // Q3 release window: one-frame runs, straight to the walk
        float e0, E = 0.0f;
        int m;
        if (st == kSustain || st == kIdle) {
            e0 = st == kIdle ? 0.0f : h.sustain;
            if (st != kIdle) h.env = h.sustain;
            m = runEnd - i;
        } else {
            e0 = adsr_next(h);                                     // exact step; may clamp and change state
            m = 1;
            const int X = runEnd - i - 1;
            float step = st == kAttack ? h.aRate : (st == kDecay ? h.dRate : h.rRate), top, bottom;
            if (h.adsrState == st && X > 0 && binade_step_f(e0, step, E, top, bottom)) {
                int x;
                if (st == kAttack) x = env_run_up(e0, E, fminf(1.0f, top), X);
                else {
                    // stay strictly above the binade floor (a result below it would round on the finer grid) and,
                    // decaying, strictly above sustain (<= sustain ends the decay, juce_ADSR.h:170)
                    const float floorUp = __int_as_float(__float_as_int(bottom) + 1);
                    const float lim = st == kDecay ? fmaxf(floorUp, __int_as_float(__float_as_int(h.sustain) + 1)) : floorUp;
                    x = env_run_down(e0, E, lim, X);
                    E = -E;
                }
                h.env = __fmaf_rn((float) x, E, e0);
                m = 1 + x;
                if (x == 0) E = 0.0f;
            } else E = 0.0f;
        }
        if (nseg == 0) { ep.e0a = e0; ep.Ea = E; } else { ep.e0b = e0; ep.Eb = E; ep.ce = i; }
        ++nseg;
        i += m;
        if (i == k0) { adsr_note_off(h); b.releaseArmed = 1; }
        if (!JMODE && h.adsrState == kIdle && !(pl.stopped && i == nPos)) { ended = true; return i; }
    }
    if (i >= nPos) return nPos;
    ep.ct = i;
    if (nseg == 1) ep.ce = i;
    if (nseg == 0) ep.ce = 0;
    resume = i;
    return -1;
}

// Part 2 - frame by frame from tile frame i on, into the table row te (indexed by tile frame).
template <bool JMODE, bool Q3>
__device__ __forceinline__ int plan_envelope_walk(VoiceHot& h, VoiceBlockCtl& b, const PosPlan& pl, float* te, int i, bool& ended) {
    ended = false;
    const int nPos = pl.nPos;
    const int lastNoteOffInc = pl.stopped ? nPos - 1 : nPos;
    const double relRcp = __ddiv_rn(1.0, h.relDen);
    if (!JMODE && Q3 && h.adsrState == kRelease && h.release > 0.0f && pl.kRel <= i + 1) {
        // Quirk Q3 inside the release window: every frame is  env -= releaseRate  (juce_ADSR.h:178-184) followed by a
        // noteOff that re-derives releaseRate = env / (release * sr)  (:132-146) - a serial recurrence with no closed
        // form.  Same arithmetic as the generic loop below, without its state tests; the frame that takes the
        // envelope to <= 0 (and everything after it) is left to the generic loop.
        const int iEnd = min(nPos, lastNoteOffInc);              // frames i < iEnd are followed by a noteOff
        float env = h.env, rr = h.rRate;
        // The recurrence is serial, so its latency is the planner's critical path (ncu, round 2: ~320 cycles per frame through
        // f32 -> f64 -> DMUL -> f32 and two data-dependent branches).  Fast form, all f32, 20 cycles of dependent issue: with
        // D = relDen split as Dh + Dl and r = RN32(1 / D), q0 = RN(e r) is within 2 ulp of x = e / D, res = e - q0 D is taken
        // by two FMAs, and q0 + res r is evaluated with r widened and narrowed by 2^-14: both roundings bracket x (and
        // RN64(x)), so where they agree they ARE RN32(RN64(e / D)), the reference's value; where they differ (6e-5 of the
        // frames: x within 2^-15 ulp of a rounding boundary) or e is tiny, the block falls back to the f64 form below.
        // Checked against (float) ((double) e / D) on 1.5e9 random (e, D): no mismatch.
        const float Dh = __double2float_rn(h.relDen), Dl = __double2float_rn(__dsub_rn(h.relDen, (double) Dh));
        const float rf = __double2float_rn(relRcp);
        const float ra = __fmul_rn(rf, 1.00006103515625f), rb = __fmul_rn(rf, 0.99993896484375f);
        const bool fastOk = h.relDen >= 1.0 && h.relDen < 1.0e9;   // the range the f32 form was checked on
        for (;;) {
            if (fastOk && i + 8 <= iEnd) {
                float ee[8];
                float en = env, rn = rr;
                bool bad = false;
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const float e = __fsub_rn(en, rn);
                    const float q0 = __fmul_rn(e, rf);
                    const float r1 = __fmaf_rn(-q0, Dh, e);
                    const float r2 = __fmaf_rn(-q0, Dl, r1);
                    const float qa = __fmaf_rn(r2, ra, q0), qb = __fmaf_rn(r2, rb, q0);
                    bad = bad || !(e > 1.0e-30f) || qa != qb;
                    ee[k] = e; en = e; rn = qa;
                }
                if (!bad) {
                    // each lane writes its own row: two 16-byte stores cost the LSU a quarter of eight 4-byte ones
                    if ((i & 3) == 0) {
                        *reinterpret_cast<float4*>(te + i) = make_float4(ee[0], ee[1], ee[2], ee[3]);
                        *reinterpret_cast<float4*>(te + i + 4) = make_float4(ee[4], ee[5], ee[6], ee[7]);
                    } else {
#pragma unroll
                        for (int k = 0; k < 8; ++k) te[i + k] = ee[k];
                    }
                    env = en; rr = rn; i += 8;
                    continue;
                }
            }
            const int stop = min(iEnd, i + 8);                   // the same frames by the f64 form, which also ends the run
            bool term = false;
            while (i < stop) {
                const float e = __fsub_rn(env, rr);
                if (!(e > 0.0f)) { term = true; break; }
                te[i] = e; env = e; ++i;
                const double q = __dmul_rn((double) e, relRcp);
                const unsigned lo = (unsigned) __double2loint(q) & 0x1fffffffu;
                rr = (lo - 0x0ffffff8u <= 16u || !(q > 1.0e-30)) ? __double2float_rn(__ddiv_rn((double) e, h.relDen)) : __double2float_rn(q);
            }
            if (term || i >= iEnd) break;
        }
        h.env = env; h.rRate = rr; b.releaseArmed = 1;
    }
    for (; i < nPos; ++i) {
        te[i] = adsr_next(h);
        const int k = i + 1;
        if (!JMODE && k >= pl.kRel && k <= lastNoteOffInc) {
            if (Q3 || !b.releaseArmed) { adsr_note_off_fast(h, relRcp); b.releaseArmed = 1; }
        }
        if (!JMODE && h.adsrState == kIdle && !(pl.stopped && k == nPos)) { ended = true; return k; }
    }
    return nPos;
}

}  // namespace zl
