// zl_render.cu — the batched sampler render kernels (sm_100a).
//
// One JACK period (block) of every session = two kernels on one stream:
//
//   plan_kernel    thread = voice.  SamplerChannel::process's ring drain + handleCommand (lib/SamplerSynth.cpp:118-122,
//                  187-230), the per-block prelude of SamplerSynthVoice::process (lib/SamplerSynthVoice.cpp:176-196) and
//                  the voice's state machine for the whole block, advanced tile by tile (kT = 128 frames) in O(1) per
//                  tile: the f64 position recurrence p += ratio is exact-linear inside a binade (zl_voice.cuh), so a tile
//                  is at most two (p0, R) segments broken at binade crossings / loop wraps / the stop position; the
//                  juce::ADSR envelope is up to two exact-linear f32 segments and is stepped frame by frame into a table
//                  row only where no closed form exists (attack from 0, quirk Q3's re-armed release); positions that
//                  have no closed form (p < 1, rounding ties, tiny loops) are walked into a position row.  Output: one
//                  64-byte TileDesc per voice-tile + one 32-byte VoiceGather per voice-block (L2-resident hand-over).
//   gather_kernel  warp = bus (SamplerChannel), lane = frame.  Pure streaming: for each tile it walks the bus's voices in
//                  slot order (the serial += order of the reference), evaluates p = p0 + k*R with one DFMA, splits it
//                  into (pos, alpha), gathers both taps of four 32-frame chunks from the HBM-resident interleaved clip
//                  (lanes read consecutive source frames: coalesced, 8 x LDG.64 in flight per lane), applies gain /
//                  envelope / M-S pan (lib/SamplerSynthVoice.cpp:198-221, J: juce_Sampler.cpp:128-152), keeps the bus
//                  accumulators in registers and runs JackPassthroughPrivate::process (lib/JackPassthrough.cpp:45-115)
//                  and the AudioLevels integer peak scan (lib/AudioLevels.cpp:357-385) on them; 128-byte coalesced
//                  stores.  No shared memory, no barriers; small enough to stay in the instruction cache and to run at
//                  high occupancy, which is what hides the HBM latency of the gather.
//   master_mix_kernel  the JACK port mix of each session's dry outputs (the only cross-bus step), once per launch.
#include <cstdio>
#include <cstdlib>

#include "zl_voice.cuh"

namespace zl {

#ifndef ZL_KT
#define ZL_KT 128
#endif
// gather_kernel CTA = ONE warp = one bus: buses differ in how many voices sound, and a warp that is done hands its
// registers and shared memory to the next bus at once instead of waiting for three CTA-mates (4-warp CTAs: 0.75 ms per
// launch, 1-warp CTAs at the same 20 warps per SM: 0.65 ms; sized for 21 per SM ptxas settles on 78 registers and 22
// fit beside their 9 KB of shared memory: 0.60 ms).
#ifndef ZL_GATHER_THREADS
#define ZL_GATHER_THREADS 32
#endif
#ifndef ZL_GATHER_MINB
#define ZL_GATHER_MINB 28
#endif
#define ZL_LAUNCH(kernel, grid, block, stream, ...) kernel<<<grid, block, 0, stream>>>(__VA_ARGS__)

// Diagnostic build only (-DZL_CTA_TRACE=1, never shipped): every CTA of gather_kernel of period 8 and of plan_kernel /
// walk_kernel of period 9 (which run beside it) leaves (SM id, clock64 at entry, clock64 at exit); render_dump_trace writes the
// table of the last render to a file.  profiles/r02/cta_trace.py turns it into resident-CTA counts per SM over time.
#ifndef ZL_CTA_TRACE
#define ZL_CTA_TRACE 0
#endif
#if ZL_CTA_TRACE
constexpr int kTraceMax = 70000;
__device__ unsigned long long g_trace[3][kTraceMax][3];
__device__ __forceinline__ unsigned trace_smid() { unsigned r; asm volatile("mov.u32 %0, %%smid;" : "=r"(r)); return r; }
#define ZL_TRACE_BEGIN(cond) const bool trOn = (cond); long long trT0 = 0; if (trOn) trT0 = clock64();
#define ZL_TRACE_END(kind) if (trOn && (threadIdx.x == 0) && blockIdx.x < kTraceMax) { g_trace[kind][blockIdx.x][0] = trace_smid(); g_trace[kind][blockIdx.x][1] = (unsigned long long) trT0; g_trace[kind][blockIdx.x][2] = (unsigned long long) clock64(); }
#else
#define ZL_TRACE_BEGIN(cond)
#define ZL_TRACE_END(kind)
#endif

// Diagnostic build only (-DZL_BOUNDS_CHECK=1, never shipped; compute-sanitizer is closed on this pool): every address the render
// kernels form from planner data is tested against the object it must stay in - source taps against the clip's length + 8
// frames, staged windows against the clip and the slot, table rows against the tile, queue slots against the queue, output
// indices against the period - and violations are counted per site; render_dump_trace prints the counters.
#ifndef ZL_BOUNDS_CHECK
#define ZL_BOUNDS_CHECK 0
#endif
#if ZL_BOUNDS_CHECK
__device__ unsigned long long g_bounds[16];                   // [0] checks made, [k] violations at site k
#define ZL_CHECK(site, cond) do { atomicAdd(&g_bounds[0], 1ull); if (!(cond)) atomicAdd(&g_bounds[site], 1ull); } while (0)
#else
#define ZL_CHECK(site, cond)
#endif

constexpr int kT = ZL_KT;            // frames per tile (multiple of 128)
constexpr int kNC = kT / 32;         // 32-frame chunks per tile
constexpr int kG = 4;                // chunks gathered together (all their loads in flight at once); a tile is kNG such groups
constexpr int kNG = kNC / kG;
static_assert(kT % (32 * kG) == 0, "a tile is a whole number of 128-frame groups");

constexpr int kSlotBit = 1 << 30;     // positions come from the voice-tile's position row
constexpr int kSimpleBit = 1 << 29;   // one position segment, one envelope segment, stereo, every tap inside the sample
constexpr int kRingBit = 1 << 28;     // simple, full tile, and the source window fits one staging slot: p0b = window address, Rb = (first frame, bytes)
constexpr int kCountMask = (1 << 20) - 1;
constexpr int kSlotBytes = 2560 * (kT / 128);   // one staged source window: kT frames x ratio 2.4 x 8 B + halo, in 512-byte rows
#ifndef ZL_SLOTS
#define ZL_SLOTS 2
#endif
constexpr int kSlots = ZL_SLOTS;             // staging slots per warp: kSlots - 1 windows in flight behind the one being gathered
// Window staging engine: 1 = TMA bulk copy (cp.async.bulk, SASS UBLKCP) issued by one lane and landing on an mbarrier per
// slot; 0 = cp.async 16 B per lane (SASS LDGSTS.128, up to five per lane and window, each preceded by three predicated-off
// LDS that ptxas inserts on sm_100a).
#ifndef ZL_TMA
#define ZL_TMA 1
#endif
// Experiment switch: L2 prefetch of a tile's later windows at tile start (0 off, 1 bulk prefetch per window, 2 per-line prefetch).
#ifndef ZL_L2PF
#define ZL_L2PF 0
#endif
// The window queue runs across tile boundaries: the next tile's first window is staged while the last one of this tile is gathered.
#ifndef ZL_PRESTAGE
#define ZL_PRESTAGE 1
#endif
constexpr int kWantPeak = 1 << 8;     // VoiceGather.nch flag: the positions model takes this block's peak  :265-267

// The gather of one voice x one tile is split in two so all of the tile's loads are in flight together:
// gather_issue computes (pos, alpha, env) per lane-frame and issues both tap loads of every chunk; gather_finish does
// the arithmetic.  Lanes past the voice's last frame read zeros (from the clip's zero padding or not at all), so they add
// +0 to the bus and to the voice peak: no validity predicate is needed downstream.
struct Taps {                        // one group of kG chunks
    float2 a[kG], b[kG];             // taps at pos and pos + 1 (mono: .y duplicates .x, so r == l falls out of the same math)
    float alpha[kG], env[kG];
};

// (int) p and p - (int) p without int<->f64 conversions: p + 2^52 rounded down is floor(p) + 2^52
__device__ __forceinline__ int split_pos(double p, float& alpha) {
    const double fl = __dadd_rd(p, 4503599627370496.0);
    alpha = __double2float_rn(__dsub_rn(p, __dsub_rn(fl, 4503599627370496.0)));
    return __double2loint(fl);
}

template <bool JMODE>
__device__ __forceinline__ void gather_issue(const TileDesc& d, const VoiceGather& g, const float* __restrict__ terow,
                                             const double* __restrict__ psrow, int lane, int nchunks, int c0, Taps& t) {
    const int nrend = d.nrend & kCountMask;
    const bool slotPos = (d.nrend & kSlotBit) != 0;
    const int nch = g.nch & 0xff;
#pragma unroll
    for (int ch = 0; ch < kG; ++ch) {
        t.a[ch] = make_float2(0.0f, 0.0f); t.b[ch] = t.a[ch]; t.alpha[ch] = 0.0f; t.env[ch] = 0.0f;
        if (ch < nchunks) {                                           // warp-uniform (nchunks: chunks of this group that hold frames)
            const int c = c0 + ch * 32 + lane;
            const bool v = c < nrend;
            double p;
            ZL_CHECK(1, c >= 0 && c < kT);
            if (slotPos) p = v ? psrow[c] : 0.0;
            else {
                const bool sb = c >= d.cb;
                p = __fma_rn((double) (c - (sb ? d.cb : 0)), sb ? d.Rb : d.Ra, sb ? d.p0b : d.p0a);
            }
            const int pos = split_pos(p, t.alpha[ch]);
            float env;
            if (c >= d.ct) env = v ? terow[c] : 0.0f;
            else {
                const bool sb = c >= d.ce;
                env = __fmaf_rn((float) (c - (sb ? d.ce : 0)), sb ? d.Eb : d.Ea, sb ? d.e0b : d.e0a);   // exact: see zl_voice.cuh
            }
            t.env[ch] = env;
            const bool inR = v && (JMODE || g.sampleDuration > pos);      // Z: taps outside the sample read as 0  (:204-205)
            if (inR) {
                ZL_CHECK(2, pos >= 0 && pos + 1 <= g.sampleDuration + 8);   // frames 0 .. length + 7 exist (sampleDuration = length - 1)
                if (nch > 1) {
                    t.a[ch] = __ldg(reinterpret_cast<const float2*>(g.data) + pos);
                    t.b[ch] = __ldg(reinterpret_cast<const float2*>(g.data) + pos + 1);
                } else {
                    const float a = __ldg(g.data + pos), b = __ldg(g.data + pos + 1);
                    t.a[ch] = make_float2(a, a); t.b[ch] = make_float2(b, b);
                }
            }
        }
    }
}

// The common case (kSimpleBit): one position segment, one envelope segment, stereo, every tap inside the sample.
// FULL: all kT frames are rendered -> no range test at all.  Otherwise lanes past nrend are pointed at the clip's zero
// padding (frames length .. length + 7 are zero), so the loads stay unpredicated.
template <bool FULL>
__device__ __forceinline__ void gather_issue_simple(const TileDesc& d, const VoiceGather& g, int lane, double laneD, float laneF, int c0, Taps& t) {
    const int nrend = d.nrend & kCountMask;
    const float2* __restrict__ src = reinterpret_cast<const float2*>(g.data);
    const int zeroPos = g.sampleDuration + 2;                         // length + 1 (and + 2): inside the padding
#pragma unroll
    for (int ch = 0; ch < kG; ++ch) {
        const double p = __fma_rn(__dadd_rn(laneD, (double) (c0 + ch * 32)), d.Ra, d.p0a);
        int pos = split_pos(p, t.alpha[ch]);
        if (!FULL) pos = (c0 + ch * 32 + lane < nrend) ? pos : zeroPos;
        ZL_CHECK(3, pos >= 0 && pos + 1 <= g.sampleDuration + 8);
        t.a[ch] = __ldg(src + pos); t.b[ch] = __ldg(src + pos + 1);
    }
    if (d.Ea == 0.0f) {                                               // warp-uniform (sustain, J-mode idle)
#pragma unroll
        for (int ch = 0; ch < kG; ++ch) t.env[ch] = d.e0a;
    } else {
#pragma unroll
        for (int ch = 0; ch < kG; ++ch) t.env[ch] = __fmaf_rn(__fadd_rn(laneF, (float) (c0 + ch * 32)), d.Ea, d.e0a);
    }
}

// Interpolation, gains, (Z) M/S pan, bus accumulate.  J: juce_Sampler.cpp:128-152, Z: lib/SamplerSynthVoice.cpp:198-221
// The left and right channels go through the same chain, so they ride together in Blackwell's packed FP32x2 instructions
// (FMUL2 / FADD2: two independently IEEE-rounded f32 operations per lane per issue slot; no contraction, _rn intrinsics).
// acc[ch] = (L, R) of the bus for this lane's frame of chunk ch.
template <bool JMODE, bool Q1, bool FULL>
__device__ __forceinline__ float gather_finish(const VoiceGather& g, const Taps& t, int nchunks, float nz, float2 pm1, float2* __restrict__ acc) {
    float vmax = 0.0f;
    const float2 g2 = make_float2(g.gain, g.gain), vol2 = make_float2(g.vol, g.vol), nz2 = make_float2(nz, nz);
    const float2 pan2 = make_float2(g.lPan, g.rPan);                  // pm1 = (1, -1), opaque like nz (kept in a uniform register pair)
#pragma unroll
    for (int ch = 0; ch < kG; ++ch) {
        if (FULL || ch < nchunks) {                                   // warp-uniform
            const float alpha = t.alpha[ch], env = t.env[ch];
            const float ia = __fsub_rn(1.0f, alpha);
            const float2 al2 = make_float2(alpha, alpha), ia2 = make_float2(ia, ia), env2 = make_float2(env, env);
            // ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 (it does not for the scalar forms), which would change
            // the rounding: every product that feeds an add is therefore written as fma(x, y, nz) with nz = -0.0f that the
            // compiler cannot see through (derived from a kernel parameter) - exactly RN(x*y), sign of zero included.
            float2 lr;
            if (JMODE) {
                lr = __fadd2_rn(__ffma2_rn(t.a[ch], ia2, nz2), __ffma2_rn(t.b[ch], al2, nz2));
                const float ge = __fmul_rn(g.gain, env);
                lr = __ffma2_rn(lr, make_float2(ge, ge), nz2);
            } else {
                if (Q1)   // lib/SamplerSynthVoice.cpp:204-205 as written: gain/env/volume multiply only the second tap
                    lr = __fadd2_rn(__ffma2_rn(t.a[ch], ia2, nz2),
                                    __ffma2_rn(__fmul2_rn(__fmul2_rn(__fmul2_rn(t.b[ch], al2), g2), env2), vol2, nz2));
                else
                    lr = __fmul2_rn(__fmul2_rn(__fmul2_rn(__fadd2_rn(__ffma2_rn(t.a[ch], ia2, nz2), __ffma2_rn(t.b[ch], al2, nz2)), g2), env2), vol2);
                // M/S pan  :208-216:  m = 0.5 (l + r), s = l - r, l' = lPan m + s, r' = rPan m - s, every operation rounded on
                // its own.  Packed: (l + r, l - r) = fma((1, -1), r, l) and (l', r') = fma((1, -1), s, (lPan m, rPan m)) - a
                // product with +-1 is exact, so each half is exactly the rounded sum / difference (zero signs included: the
                // product -1 * s carries the sign a subtraction of s would see).
                const float2 ss = __ffma2_rn(pm1, make_float2(lr.y, lr.y), make_float2(lr.x, lr.x));
                const float m = __fmul_rn(0.5f, ss.x);                    // 0.5*(l+r) in f64 is exact
                lr = __ffma2_rn(pm1, make_float2(ss.y, ss.y), __fmul2_rn(pan2, make_float2(m, m)));
            }
            acc[ch] = __fadd2_rn(acc[ch], lr);                      // lanes past the voice's end add +0
            vmax = fmaxf(vmax, __fadd_rn(lr.x, lr.y));
        }
    }
    return vmax;
}

// JackPassthroughPrivate::process for one output pair   lib/JackPassthrough.cpp:58-111
__device__ __forceinline__ void passthrough_pair(const zl_passthrough_params& p, float amt, float inL, float inR,
                                                 float& oL, float& oR) {
    if (p.muted) { oL = 0.0f; oR = 0.0f; return; }
    if (p.pan_amount == 0.0f && amt == 0.0f) { oL = 0.0f; oR = 0.0f; return; }
    if (p.pan_amount == 0.0f && amt == 1.0f) { oL = inL; oR = inR; return; }
    oL = __fmul_rn(__fmul_rn(amt, inL), fminf(__fsub_rn(1.0f, p.pan_amount), 1.0f));
    oR = __fmul_rn(__fmul_rn(amt, inR), fminf(__fadd_rn(1.0f, p.pan_amount), 1.0f));
}

__device__ __forceinline__ int peak_int(float x) {   // lib/AudioLevels.cpp:365-371
    return __float2int_rz(fabsf(__fmul_rn(131072.0f, x)));
}

// Simple full tiles whose source window [first tap, last tap + 1] fits a staging slot are gathered through shared memory:
// the planner leaves the window's 16-byte aligned address and size in the (unused) second position segment.
__device__ __forceinline__ void mark_ring(TileDesc& d, const float* data, int sampleDur, int cnt, int lastPos) {
    if (cnt != kT || !(d.nrend & kSimpleBit)) return;
    const int base = (int) d.p0a & ~1;                                   // 16-byte aligned first frame (stereo: 8 B per frame)
    const int bytes = ((lastPos + 2 - base) * 8 + 511) & ~511;           // frames base .. lastPos + 1, in whole 512-byte rows
    // the rows must fit a slot and stay inside the clip buffer (length + 8 frames, sampleDur = length - 1)
    if (bytes > kSlotBytes || base * 8 + bytes > (sampleDur + 9) * 8 || (reinterpret_cast<uintptr_t>(data) & 15)) return;
    d.nrend |= kRingBit;
    d.p0b = __longlong_as_double((long long) (reinterpret_cast<uintptr_t>(data) + (size_t) base * 8));
    d.Rb = __hiloint2double(base, bytes);
    // The tile's first position relative to the window, p0a - base, in the two unused floats of envelope segment b
    // (a ring tile has one envelope segment).  Exact: both are multiples of the binade's ulp u and the difference is < 2^9,
    // so it needs no more than 53 bits; the same holds for every p_k - base of the tile, and frac(p_k - base) = frac(p_k).
    const double prel = __dsub_rn(d.p0a, (double) base);
    d.e0b = __int_as_float(__double2loint(prel));
    d.Eb = __int_as_float(__double2hiint(prel));
}

// What a position plan may change before it is accepted.
struct PlanUndo {
    double p; uint64_t nextLoopTick, nextLoopUsecs; int32_t fWrap, spanStart, lastPos; uint32_t span;
};
__device__ __forceinline__ void plan_save(PlanUndo& u, const VoiceHot& h, const VoiceBlockCtl& b) {
    u.p = h.p; u.nextLoopTick = h.nextLoopTick; u.nextLoopUsecs = h.nextLoopUsecs;
    u.fWrap = b.fWrap; u.spanStart = b.spanStart; u.lastPos = b.lastPos; u.span = b.span;
}
__device__ __forceinline__ void plan_restore(const PlanUndo& u, VoiceHot& h, VoiceBlockCtl& b) {
    h.p = u.p; h.nextLoopTick = u.nextLoopTick; h.nextLoopUsecs = u.nextLoopUsecs;
    b.fWrap = u.fWrap; b.spanStart = u.spanStart; b.lastPos = u.lastPos; b.span = u.span;
}

// ---------------------------------------------------------------------------------------------------------------------
// One voice, one block: prelude, tiles, epilogue.  Returns non-zero (and changes nothing) when DEFER is set and the voice
// is not on the event-free fast path; such voices are queued and planned by walk_kernel, packed 32 to a warp, instead
// of holding a whole warp of plan_kernel hostage with a few busy lanes.
template <bool JMODE, bool Q3, bool DEFER>
__device__ __forceinline__ int plan_voice_block(const RenderParams& P, int vIdx, int gb, unsigned long long& nVoiceFrames,
                                                 unsigned long long& nSrcFloats) {
    const int blk = P.blk, nframes = P.nframes;
    VoiceDev* rec = P.voices + vIdx;
    VoiceHot h;
    hot_load(h, *rec);
    const zl_block_clock ck = P.clocks[blk];
    // ---- block prelude ---------------------------------------------------------------------------------------
    VoiceBlockCtl bc;
    bc.stopD = 0.0; bc.startD = 0.0; bc.relStart = 0.0; bc.usPerFrame = 0.0; bc.lengthInTicks = 0; bc.fWrap = nframes;
    bc.beatAligned = 0; bc.spanStart = 0; bc.lastPos = 0; bc.span = 0; bc.nch = 0; bc.clipFrames = 0; bc.releaseArmed = 0;
    VoiceGather g;
    g.data = nullptr; g.sampleDuration = 0; g.nch = 0; g.gain = g.vol = g.lPan = g.rPan = 0.0f;
    bool active = false;
    if (P.buses[gb].enabled && (h.flags & VF_PLAYING))                              // :136-140
        active = voice_block_begin<JMODE>(h, P, ck, nframes, bc, g);
    const bool walked = active;
    const int sampleDuration = g.sampleDuration;
    int nrendBlock = 0;

    // ---- fast path: the whole block is one event-free run --------------------------------------------------
    // Sustaining voice whose positions stay inside one binade and below every event of :223-261 (loop wrap, stop
    // position, release window) until the end of the block: every tile is one (p0, R) segment with a constant
    // envelope, exactly what the generic planner below would produce tile by tile.
    bool fast = false;
    // The envelope of the whole period as ONE exact-linear run e(c) = e0 + c E (zl_voice.cuh): sustain (E = 0), or an attack
    // under way / a decay / a plain release that stays inside one f32 binade and short of its next state until the period
    // ends.  The generic planner would find the same run tile by tile (each tile's first step RN(e + rate) lands on the same
    // grid), so the descriptors - and the voice's state after the period - are identical.
    float envE0 = h.sustain, envE = 0.0f, envLast = h.sustain;
    bool envOk = h.adsrState == kSustain;
    if (active && !envOk && nframes > 1) {
        const int st = h.adsrState;
        if (JMODE && st == kIdle) { envOk = true; envE0 = 0.0f; envLast = h.env; }
        else if ((st == kAttack && h.env > 0.0f) || st == kDecay || st == kRelease) {
            VoiceHot hc = h;
            const float e0 = adsr_next(hc);                        // exact step; may clamp and change state
            const float step = st == kAttack ? h.aRate : (st == kDecay ? h.dRate : h.rRate);
            float E, topF, bottomF;
            if (hc.adsrState == st && binade_step_f(e0, step, E, topF, bottomF)) {
                const int X = nframes - 1;
                int x;
                if (st == kAttack) x = env_run_up(e0, E, fminf(1.0f, topF), X);
                else {
                    const float floorUp = __int_as_float(__float_as_int(bottomF) + 1);
                    const float lim = st == kDecay ? fmaxf(floorUp, __int_as_float(__float_as_int(h.sustain) + 1)) : floorUp;
                    x = env_run_down(e0, E, lim, X);
                    E = -E;
                }
                if (x == X) { envOk = true; envE0 = e0; envE = E; envLast = __fmaf_rn((float) X, E, e0); }
            }
        }
    }
    if (active && envOk) {
        double R, top;
        if (binade_step(h.p, h.ratio, R, top)) {
            const double q = __fma_rn((double) nframes, R, h.p);   // position after the block's last increment
            if (!JMODE && h.looping) fast = bc.beatAligned ? (bc.fWrap >= nframes && q < top) : (q < fmin(top, bc.stopD));
            else fast = q < fmin(top, bc.stopD) && (JMODE || q < bc.relStart);
            if (fast) {
                // Tile t starts at p_(t kT) = h.p + t kT R and e_(t kT) = envE0 + t kT envE.  Both runs stay on one grid for the
                // whole period, so every one of these values is representable and the running sums below ARE them: the
                // same bits as fma(t0, R, h.p) / fmaf(t0, envE, envE0), without an int -> f64 conversion per tile.
                double pT = h.p;
                float eT = envE0;
                const double RT = __dmul_rn(R, (double) kT);
                const float ET = __fmul_rn(envE, (float) kT);
                const int simple = bc.nch == 2 ? kSimpleBit : 0;
                TileDesc* dst = P.descs + vIdx;
                int left = nframes;
                for (int t = 0; t < P.tilesPerBlock; ++t) {
                    const int n = min(kT, left);
                    TileDesc d;
                    d.p0a = pT; d.Ra = R; d.p0b = 0.0; d.Rb = 0.0;
                    d.e0a = eT; d.Ea = envE; d.e0b = 0.0f; d.Eb = 0.0f;
                    d.cb = n; d.ce = n; d.ct = INT_MAX;
                    const int lastPos = (int) __fma_rn(n == kT ? (double) (kT - 1) : (double) (n - 1), R, pT);
                    d.nrend = n | ((JMODE || lastPos < sampleDuration) ? simple : 0);
                    mark_ring(d, g.data, sampleDuration, n, lastPos);
                    *dst = d;
                    bc.lastPos = lastPos;
                    dst += P.nVoices; left -= kT;
                    pT = __dadd_rn(pT, RT); eT = __fadd_rn(eT, ET);       // not used after the last tile
                }
                h.p = q; h.env = envLast;
                nrendBlock = nframes;
            } else if (q >= top) {
                // ---- one binade crossing, nothing else --------------------------------------------------------
                // The positions run into the top of their binade inside this period (a looping voice does that five or six
                // times per pass over a half-second clip: at 1024, 2048, ... frames - half of all the voices that used to leave
                // the fast path).  Frames [0, kx) are the run below the top; the increment after frame kx - 1 is the
                // reference's own RN(p + ratio), which lands on the next binade's grid; frames [kx, nframes) are the run
                // from there.  With no other event on either side (loop end / wrap, stop position, release window) this is
                // exactly what plan_positions + pos_explicit_step produce tile by tile: one segment per tile, two in the
                // tile that holds the crossing.
                const bool la = !JMODE && h.looping;
                const bool lim = la && bc.beatAligned ? bc.fWrap >= nframes : top <= bc.stopD;
                if (lim && (JMODE || la || top <= bc.relStart)) {
                    const int kx = first_k_reaching(h.p, R, top, nframes);          // 1 .. nframes
                    const double plast = __fma_rn((double) (kx - 1), R, h.p);
                    const double pn = __dadd_rn(plast, h.ratio);
                    double Rb, topb;
                    if (binade_step(pn, h.ratio, Rb, topb)) {
                        const double qb = __fma_rn((double) (nframes - kx), Rb, pn);
                        bool ok = qb < topb;
                        if (!(la && bc.beatAligned)) ok = ok && pn < bc.stopD && qb < bc.stopD;
                        if (!JMODE && !la) ok = ok && pn < bc.relStart && qb < bc.relStart;
                        if (ok) {
                            fast = true;
                            for (int t = 0; t < P.tilesPerBlock; ++t) {
                                const int t0 = t * kT;
                                const int n = min(kT, nframes - t0);
                                TileDesc d;
                                d.e0a = __fmaf_rn((float) t0, envE, envE0); d.Ea = envE; d.e0b = 0.0f; d.Eb = 0.0f;
                                d.ce = n; d.ct = INT_MAX; d.cb = n; d.p0b = 0.0; d.Rb = 0.0;
                                const int last = t0 + n - 1;
                                const int lastPos = (int) (last < kx ? __fma_rn((double) last, R, h.p) : __fma_rn((double) (last - kx), Rb, pn));
                                if (t0 + n <= kx || t0 >= kx) {
                                    if (t0 >= kx) { d.p0a = __fma_rn((double) (t0 - kx), Rb, pn); d.Ra = Rb; }
                                    else { d.p0a = __fma_rn((double) t0, R, h.p); d.Ra = R; }
                                    d.nrend = n | ((bc.nch == 2 && (JMODE || lastPos < sampleDuration)) ? kSimpleBit : 0);
                                    mark_ring(d, g.data, sampleDuration, n, lastPos);
                                } else {
                                    d.p0a = __fma_rn((double) t0, R, h.p); d.Ra = R; d.cb = kx - t0; d.p0b = pn; d.Rb = Rb;
                                    d.nrend = n;
                                }
                                P.descs[(size_t) t * P.nVoices + vIdx] = d;
                                bc.lastPos = lastPos;
                            }
                            h.p = qb; h.env = envLast;
                            nrendBlock = nframes;
                        }
                    }
                }
            }
        }
    }

    // plan_kernel (DEFER) takes only the event-free voices: everything else - envelope segments in attack/decay/release,
    // loop wraps, binade crossings, the frame-by-frame walks of an attack from 0 and of quirk Q3's release window - would
    // drag the other 31 lanes of its warp through the generic planner, so it is queued for walk_kernel, which runs the
    // same code with those voices packed 32 to a warp.  Nothing has been written yet.
    if (DEFER && active && !fast && (P.dbgFlags & 8)) {          // ZL_B200_WALK_PROF: why the fast path was not taken
        int why = 0;                                             // 0 envelope, 1 binade, 2 beat-aligned wrap, 3 loop end/top, 4 one-shot stop/top, 5 release window
        if (envOk) {
            double R, top;
            if (!binade_step(h.p, h.ratio, R, top)) why = 1;
            else {
                const double q = __fma_rn((double) nframes, R, h.p);
                if (!JMODE && h.looping) why = bc.beatAligned ? 2 : 3;
                else why = !(q < fmin(top, bc.stopD)) ? 4 : 5;
                if (why == 3 && !(q < bc.stopD)) why = 6;        // 6: loop end proper (3: binade top)
            }
        }
        atomicAdd(P.voiceFrames + 24 + why, 1ull);
    }
    if (DEFER && active && !fast) {
        // kind 1: will step an envelope frame by frame for (most of) the period - quirk Q3's release window of a one-shot
        // (conservative test on p reaching relStart, :253) or an attack from 0; kind 2: everything else (O(1) per tile).
        // The two kinds are queued apart so a warp of walk_kernel is not a mix of 10 us and 100 us voices.
        bool serial = h.adsrState == kAttack && h.env == 0.0f;
        if (!JMODE && Q3 && !h.looping) serial = serial || __fma_rn((double) nframes * 1.0000001, h.ratio, h.p) >= bc.relStart;
        return serial ? 1 : 2;
    }

    if (DEFER && !fast) {                                     // not playing: empty tiles (the gather skips them)
        TileDesc d;
        d.p0a = 0.0; d.Ra = 0.0; d.p0b = 0.0; d.Rb = 0.0; d.e0a = 0.0f; d.Ea = 0.0f; d.e0b = 0.0f; d.Eb = 0.0f;
        d.cb = 0; d.ce = 0; d.ct = INT_MAX; d.nrend = 0;
        for (int t = 0; t < P.tilesPerBlock; ++t) P.descs[(size_t) t * P.nVoices + vIdx] = d;
    }
    for (int t = 0; t < ((DEFER || fast) ? 0 : P.tilesPerBlock); ++t) {
        const int t0 = t * kT;
        const int n = min(kT, nframes - t0);
        const size_t vt = (size_t) t * P.nVoices + vIdx;    // voice-tile
        TileDesc d;
        d.p0a = 0.0; d.Ra = 0.0; d.p0b = 0.0; d.Rb = 0.0; d.e0a = 0.0f; d.Ea = 0.0f; d.e0b = 0.0f; d.Eb = 0.0f;
        d.cb = n; d.ce = n; d.ct = INT_MAX; d.nrend = 0;
        if (active) {
            double* ps = P.ps + vt * kT;
            PosPlan pl;
            PlanUndo undo;
            plan_save(undo, h, bc);
            plan_positions<JMODE>(h, bc, ck, t0, n, nframes, pl);
            if (pl.irregular) {
                plan_restore(undo, h, bc);
                walk_positions<JMODE>(h, bc, ck, t0, n, nframes, ps, pl);
            }
            EnvPlan ep;
            bool ended;
            int resume;
            int cnt = plan_envelope_closed<JMODE, Q3>(h, bc, pl, ep, ended, resume);
            if (resume >= 0) cnt = plan_envelope_walk<JMODE, Q3>(h, bc, pl, P.te + vt * kT, resume, ended);
            d.p0a = pl.p0a; d.Ra = pl.Ra; d.p0b = pl.p0b; d.Rb = pl.Rb; d.cb = pl.cb;
            d.e0a = ep.e0a; d.Ea = ep.Ea; d.e0b = ep.e0b; d.Eb = ep.Eb; d.ce = ep.ce; d.ct = ep.ct;
            d.nrend = cnt | (pl.irregular ? kSlotBit : 0);
            if (cnt > 0 && !pl.irregular && pl.cb >= cnt && ep.ce >= cnt && ep.ct >= cnt && bc.nch == 2) {
                const int lastPos = (int) __fma_rn((double) (cnt - 1), pl.Ra, pl.p0a);
                if (JMODE || lastPos < sampleDuration) { d.nrend |= kSimpleBit; mark_ring(d, g.data, sampleDuration, cnt, lastPos); }
            }
            if (ended && cnt < pl.nPos) {                          // the ADSR ran out before the tile did: p is frame cnt's
                if (pl.irregular) h.p = ps[cnt];
                else h.p = cnt >= pl.cb ? __fma_rn((double) (cnt - pl.cb), pl.Rb, pl.p0b) : __fma_rn((double) cnt, pl.Ra, pl.p0a);
            }
            if (pl.stopped || ended) { voice_stop(h, false); active = false; }
            nrendBlock += cnt;
        }
        P.descs[vt] = d;
    }

    // ---- block epilogue ----------------------------------------------------------------------------------
    if (walked) {
        nVoiceFrames = (unsigned long long) nrendBlock;
        if (nrendBlock > 0) nSrcFloats = (unsigned long long) (bc.span + (uint32_t) (bc.lastPos + 2 - bc.spanStart)) * (unsigned) bc.nch;
        // positions model update   lib/SamplerSynthVoice.cpp:265-267.  Z: the gain is this block's peak of l + r, which only
        // the gather knows: it is reset here and max-ed in by gather_kernel (kWantPeak).
        if (h.clip >= 0 && (JMODE || (h.flags & VF_HAS_POSID))) {
            // J: gain 0.  Z, V <= 32: gather_kernel stores the peak (a plain store by the lane that owns the voice, so this
            // kernel may already be planning the next period on another stream).  Z, V > 32: reset here, atomicMax there.
            if (JMODE || P.V > 32) rec->posGain = 0.0f;
            if (!JMODE) g.nch |= kWantPeak;
            h.posProgress = __double2float_rn(__ddiv_rn(h.p, h.srcLen));
        }
    }
    P.vgs[vIdx] = g;
    hot_store(h, *rec);
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
// plan_kernel: thread = voice, CTA = whole buses (planBusesPerCta x V threads, rounded up to a warp).
#ifndef ZL_PLAN_MINB
#define ZL_PLAN_MINB 6
#endif
template <bool JMODE, bool Q3, int MAXT>
__global__ void __launch_bounds__(MAXT, MAXT == 128 ? ZL_PLAN_MINB : 1) plan_kernel(const RenderParams P) {
    const int V = P.V, B = P.B;
    const int tid = threadIdx.x;
    const int lb = tid / V, slot = tid - lb * V;              // bus within the CTA, voice slot within the bus
    const int blk = P.blk;
    ZL_TRACE_BEGIN(P.blk == 9)
    const unsigned FULL = 0xffffffffu;
    unsigned long long nVoiceFrames = 0, nSrcFloats = 0;
    // grid-stride over the bus groups: beside a running gather the grid is capped (launch_plan) so that the planner
    // shares every SM with the gather instead of displacing it
    for (int grp = blockIdx.x; grp < P.planGroups; grp += gridDim.x) {
    const int gb = grp * P.planBusesPerCta + lb;               // global bus = session * B + bus
    const bool hasVoice = lb < P.planBusesPerCta && gb < P.nSessions * B;
    const int vIdx = gb * V + slot;                            // global voice

    // ---- commands of this block: SamplerChannel::process ring drain   lib/SamplerSynth.cpp:118-122 ------------------
    // The bus's slot-0 thread applies them serially on the HBM voice records before anybody loads its state.
    if (P.cmdRange) {
        if (hasVoice && slot == 0) {
            const int c0 = P.cmdRange[2 * gb], c1 = c0 + P.cmdRange[2 * gb + 1];
            for (int c = c0; c < c1; ++c)
                if (P.cmds[c].block == blk) bus_handle_command(P.voices + (size_t) gb * V, V, gb % B - 2, P.cmds[c], P);
        }
        __syncthreads();
    }
    int kind = 0;
    if (hasVoice) {
        unsigned long long a = 0, b = 0;
        kind = plan_voice_block<JMODE, Q3, true>(P, vIdx, gb, a, b);
        if (P.dbgFlags & 4) kind = 0;
        nVoiceFrames += a; nSrcFloats += b;
    }
    // queue the deferred voices of this warp: serial walkers from the front of the queue, the others from its back; one
    // atomic per warp and kind
    const unsigned lt = (1u << (tid & 31)) - 1u;
    const unsigned mA = __ballot_sync(FULL, kind == 1), mB = __ballot_sync(FULL, kind == 2);
    if (mA | mB) {
        int baseA = 0, baseB = 0;
        if ((tid & 31) == 0) {
            if (mA) baseA = atomicAdd(P.walkCount + 2 * blk, __popc(mA));
            if (mB) baseB = atomicAdd(P.walkCount + 2 * blk + 1, __popc(mB));
        }
        baseA = __shfl_sync(FULL, baseA, 0); baseB = __shfl_sync(FULL, baseB, 0);
        ZL_CHECK(6, kind == 0 || (kind == 1 ? baseA + __popc(mA & lt) : P.nVoices - 1 - (baseB + __popc(mB & lt))) >= 0);
        ZL_CHECK(6, kind == 0 || (kind == 1 ? baseA + __popc(mA & lt) : P.nVoices - 1 - (baseB + __popc(mB & lt))) < P.nVoices);
        if (kind == 1) P.walkQueue[baseA + __popc(mA & lt)] = vIdx;
        if (kind == 2) P.walkQueue[P.nVoices - 1 - (baseB + __popc(mB & lt))] = vIdx;
    }
    }   // bus groups
    for (int o = 16; o > 0; o >>= 1) {
        nVoiceFrames += __shfl_down_sync(FULL, nVoiceFrames, o);
        nSrcFloats += __shfl_down_sync(FULL, nSrcFloats, o);
    }
    if ((tid & 31) == 0 && nVoiceFrames) {
        atomicAdd(P.voiceFrames, nVoiceFrames); atomicAdd(P.voiceFrames + 1, nSrcFloats);
        atomicAdd(P.voiceFrames + 2, nVoiceFrames); atomicAdd(P.voiceFrames + 3, nSrcFloats);
    }
    ZL_TRACE_END(1)
}


// walk_kernel: thread = queued voice (the ones plan_kernel deferred); CTA = one warp, grid-stride over the warps of
// the two queues (serial walkers first, so the long ones start first).
#ifndef ZL_WALK_MINB
#define ZL_WALK_MINB 1      // one-warp CTAs per SM the register allocation is sized for (1: no cap, 128 registers)
#endif
template <bool JMODE, bool Q3>
__global__ void __launch_bounds__(32, ZL_WALK_MINB) walk_kernel(const RenderParams P) {
    const unsigned FULL = 0xffffffffu;
    const int tid = threadIdx.x;
    ZL_TRACE_BEGIN(P.blk == 9)
    const int cA = P.walkCount[2 * P.blk], cB = P.walkCount[2 * P.blk + 1];
    const int wA = (cA + 31) >> 5, wB = (cB + 31) >> 5;
    unsigned long long nVoiceFrames = 0, nSrcFloats = 0;
    for (int w = blockIdx.x; w < wA + wB; w += gridDim.x) {
        int vIdx = -1;
        ZL_CHECK(7, cA >= 0 && cB >= 0 && cA + cB <= P.nVoices);
        if (w < wA) { const int i = w * 32 + tid; if (i < cA) vIdx = P.walkQueue[i]; }
        else { const int j = (w - wA) * 32 + tid; if (j < cB) vIdx = P.walkQueue[P.nVoices - 1 - j]; }
        // ZL_B200_WALK_PROF (profiling aid): cycles per queued voice by what it was doing when the period began -
        // [0] attack from 0, [1] release (quirk Q3's frame-by-frame window), [2] attack under way / decay, [3] sustain, and
        // [4] the whole warp - as count | sum | max in P.voiceFrames[8 + 3 k ..]
        const bool prof = (P.dbgFlags & 8) != 0;
        long long tw0 = 0;
        if (prof) tw0 = clock64();
        if (vIdx >= 0) {
            int cat = 0;
            long long t0 = 0;
            if (prof) {
                const VoiceDev& r = P.voices[vIdx];
                cat = r.adsrState == kAttack ? (r.env == 0.0f ? 0 : 2) : (r.adsrState == kRelease ? 1 : (r.adsrState == kDecay ? 2 : 3));
                t0 = clock64();
            }
            unsigned long long a = 0, b = 0;
            plan_voice_block<JMODE, Q3, false>(P, vIdx, vIdx / P.V, a, b);
            nVoiceFrames += a; nSrcFloats += b;
            if (prof) {
                const unsigned long long dt = (unsigned long long) (clock64() - t0);
                atomicAdd(P.voiceFrames + 8 + 3 * cat, 1ull); atomicAdd(P.voiceFrames + 9 + 3 * cat, dt); atomicMax(P.voiceFrames + 10 + 3 * cat, dt);
            }
        }
        if (prof) {
            __syncwarp();
            const unsigned long long dt = (unsigned long long) (clock64() - tw0);
            if (tid == 0) { atomicAdd(P.voiceFrames + 20, 1ull); atomicAdd(P.voiceFrames + 21, dt); atomicMax(P.voiceFrames + 22, dt); }
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        nVoiceFrames += __shfl_down_sync(FULL, nVoiceFrames, o);
        nSrcFloats += __shfl_down_sync(FULL, nSrcFloats, o);
    }
    if ((tid & 31) == 0 && nVoiceFrames) {
        atomicAdd(P.voiceFrames, nVoiceFrames); atomicAdd(P.voiceFrames + 1, nSrcFloats);
        atomicAdd(P.voiceFrames + 2, nVoiceFrames); atomicAdd(P.voiceFrames + 3, nSrcFloats);
    }
    ZL_TRACE_END(2)
}

// ---------------------------------------------------------------------------------------------------------------------
// gather_kernel: warp = bus, lane = frame; all tiles of one block.
__device__ __forceinline__ TileDesc load_desc(const TileDesc* p) {
    TileDesc d;
    const int4* q = reinterpret_cast<const int4*>(p);
    int4* o = reinterpret_cast<int4*>(&d);
#pragma unroll
    for (int i = 0; i < 4; ++i) o[i] = __ldg(q + i);
    return d;
}
__device__ __forceinline__ VoiceGather load_vg(const VoiceGather* p) {
    VoiceGather g;
    const int4* q = reinterpret_cast<const int4*>(p);
    int4* o = reinterpret_cast<int4*>(&g);
    o[0] = __ldg(q); o[1] = __ldg(q + 1);
    return g;
}

constexpr int kStageV = 16;          // voices per bus whose descriptors are staged through shared memory (else: straight from L2)
// Asynchronous window staging (cp.async, SASS LDGSTS.128): the warp copies the source window of voice-tile v + 1 and
// v + 2 - contiguous, 16 bytes per lane per instruction, so every 128-byte line crosses the L1 data pipe once - into
// shared-memory slots while the arithmetic of voice-tile v runs on a slot that has already landed.  No registers are
// held by loads in flight, and the taps themselves are then two LDS.64 per lane-frame.
template <int OFF> __device__ __forceinline__ void cp_async16(uint32_t saddr, const void* gptr) {   // same immediate offset on both sides
    asm volatile("cp.async.cg.shared.global [%0 + %2], [%1 + %2], 16;" :: "r"(saddr), "l"(gptr), "n"(OFF) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }

// Issues the copies of one window into the slot at shared address `slot`; the caller commits the group.
#if !ZL_TMA
__device__ __forceinline__ void stage_window(const TileDesc& d, int lane16, uint32_t slot) {
    const unsigned char* src = reinterpret_cast<const unsigned char*>((uintptr_t) __double_as_longlong(d.p0b)) + lane16;
    const int rows = __double2loint(d.Rb) >> 9;                          // whole 512-byte rows (warp-uniform), >= 1
    const uint32_t sa = slot + lane16;
    static_assert(kSlotBytes == 5 * 512, "the cp.async stage_window copies up to five rows (ZL_KT=128); other tile sizes need ZL_TMA=1");
    cp_async16<0>(sa, src);
    if (rows > 1) cp_async16<512>(sa, src);
    if (rows > 2) cp_async16<1024>(sa, src);
    if (rows > 3) cp_async16<1536>(sa, src);
    if (rows > 4) cp_async16<2048>(sa, src);
}
#else
// TMA staging: ONE lane issues ONE bulk copy per window (cp.async.bulk global -> shared, SASS UBLKCP), completion is
// counted in bytes on the slot's mbarrier (count 1: the issuing lane's arrive.expect_tx), every lane then waits for the
// slot's phase.  A slot is staged again only after the warp has consumed it (the __syncwarp behind gather_finish_ring),
// so its barrier's phase parity is just the number of times the slot went round.
__device__ __forceinline__ void mbar_init(uint32_t bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "ZL_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra ZL_DONE;\n"
        "bra ZL_WAIT;\n"
        "ZL_DONE:\n"
        "}" :: "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void stage_window_tma_raw(uint64_t src, uint32_t bytes, uint32_t slot, uint32_t bar) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "elect.sync _|p, 0xffffffff;\n"
        "@p mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n"
        "@p cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%2], [%3], %1, [%0];\n"
        "}" :: "r"(bar), "r"(bytes), "r"(slot), "l"(src) : "memory");
}
__device__ __forceinline__ void stage_window_tma(const TileDesc& d, int lane, uint32_t slot, uint32_t bar) {
    // Every lane holds the same descriptor (a shared-memory broadcast), so the operands are warp-uniform and the issuing
    // lane is picked by elect.sync inside the warp-convergent code: ptxas then feeds UBLKCP from uniform registers directly
    // instead of wrapping it in the loop it builds around a divergent `if (lane == 0)`.
    const uint64_t src = (uint64_t) __double_as_longlong(d.p0b);
    const uint32_t bytes = (uint32_t) __double2loint(d.Rb);              // whole 512-byte rows, 16-byte aligned source (mark_ring)
    (void) lane;
    ZL_CHECK(5, (src & 15) == 0 && (bytes & 15) == 0 && bytes > 0 && bytes <= (uint32_t) kSlotBytes);
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "elect.sync _|p, 0xffffffff;\n"
        "@p mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n"
        "@p cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%2], [%3], %1, [%0];\n"
        "}" :: "r"(bar), "r"(bytes), "r"(slot), "l"(src) : "memory");
}
#endif

// One group (kG chunks from tile frame c0) of a window-staged voice-tile.
// Positions run in the window's own frame: p'_k = (p0a - base) + k R (mark_ring), so floor(p'_k) indexes the slot directly
// and the fraction is that of p_k.  Every p'_k of the tile is exactly representable (zl_voice.cuh: multiples of the binade's
// ulp, one binade per segment), so the lane's first position is one DFMA and each further chunk ONE exact DADD of 32 R
// instead of a DADD + DFMA pair; the envelope e_k = e0a + k Ea is exact on its f32 grid the same way (Ea = 0: sustain).
template <bool JMODE, bool Q1>
__device__ __forceinline__ float gather_finish_ring(const TileDesc& d, const VoiceGather& g, const unsigned char* slot, double laneD, float laneF,
                                                    float nz, float2 pm1, int c0, float2* __restrict__ acc) {
    Taps t;
    const float2* win = reinterpret_cast<const float2*>(slot);                          // win[pos'] = source frame base + pos'
    const double prel = __hiloint2double(__float_as_int(d.Eb), __float_as_int(d.e0b));
    const double R32 = __dmul_rn(d.Ra, 32.0);
    const float E32 = __fmul_rn(d.Ea, 32.0f);
    double p = __fma_rn(c0 ? __dadd_rn(laneD, (double) c0) : laneD, d.Ra, prel);
    float e = __fmaf_rn(c0 ? __fadd_rn(laneF, (float) c0) : laneF, d.Ea, d.e0a);
#pragma unroll
    for (int ch = 0; ch < kG; ++ch) {
        if (ch) { p = __dadd_rn(p, R32); e = __fadd_rn(e, E32); }
        const int pos = split_pos(p, t.alpha[ch]);
        ZL_CHECK(4, pos >= 0 && (pos + 2) * 8 <= __double2loint(d.Rb) && __double2loint(d.Rb) <= kSlotBytes);
        t.a[ch] = win[pos]; t.b[ch] = win[pos + 1];
        t.env[ch] = e;
    }
    return gather_finish<JMODE, Q1, true>(g, t, kG, nz, pm1, acc);
}

// LEAN: only the raw SamplerChannel ports (and what the master mix and the meters need) are materialised - no dry / fx planes,
// no sum of squares: the per-chunk epilogue then carries none of their tests (the offline-render configuration).
template <bool JMODE, bool Q1, bool Q2, bool STAGED, bool LEAN>
__global__ void __launch_bounds__(ZL_GATHER_THREADS, ZL_GATHER_MINB) gather_kernel(const RenderParams P) {
    constexpr int QS = Q2 ? 1 : 0;                            // frame f lands at output index f + QS   :218-221
    constexpr int W = ZL_GATHER_THREADS / 32;
    // per warp: this tile's V descriptors, the block's V gather records (V <= kStageV) and the window staging slots
    __shared__ TileDesc sDesc[STAGED ? W : 1][kStageV];
    __shared__ VoiceGather sVg[STAGED ? W : 1][kStageV];
    __shared__ __align__(128) unsigned char sWin[STAGED ? W : 1][kSlots][kSlotBytes];
#if ZL_TMA
    __shared__ __align__(8) unsigned long long sBar[STAGED ? W : 1][kSlots];
#endif
    int lane = threadIdx.x & 31, warp = STAGED ? (threadIdx.x >> 5) : 0;
    asm volatile("" : "+r"(lane), "+r"(warp));            // opaque: keeps ptxas from re-deriving them from S2R inside the voice loop
#if !ZL_TMA
    const int lane16 = lane * 16;
#endif
    const uint32_t slot0 = (uint32_t) __cvta_generic_to_shared(&sWin[warp][0][0]);
    const int gbw = blockIdx.x * W + (threadIdx.x >> 5);      // global bus
    if (gbw >= P.nSessions * P.B) return;
    ZL_TRACE_BEGIN(P.blk == 8)
    const unsigned FULL = 0xffffffffu;
    const int V = P.V, nframes = P.nframes, blk = P.blk;
    const double laneD = (double) lane;
    const float laneF = (float) lane;
    const float nz = __int_as_float((int) (0x80000000u | ((unsigned) P.nframes >> 31)));   // -0.0f, opaque to the compiler (see gather_finish)
    const unsigned opq = (unsigned) P.nframes >> 31;                                       // 0, opaque
    const float2 pm1 = make_float2(__int_as_float((int) (0x3f800000u | opq)), __int_as_float((int) (0xbf800000u | opq)));   // (1, -1) for the packed M/S pan
    const zl_passthrough_params pt = P.buses[gbw].pt;
    const int64_t off = (int64_t) blk * nframes;
    const int64_t rowL = ((int64_t) gbw * 2) * P.stride + off, rowR = rowL + P.stride;
    const size_t v0 = (size_t) gbw * V;
    int pkL = 0, pkR = 0;
    double sqL = 0.0, sqR = 0.0;
    int vpk = 0;                                              // V <= 32: lane vs keeps voice vs's block peak (float bits, >= +0)
    float* const __restrict__ out0L = P.out[0] ? P.out[0] + rowL + QS : nullptr;   // LEAN: the raw rows of this bus and period
    float* const __restrict__ out0R = P.out[0] ? P.out[0] + rowR + QS : nullptr;
#if ZL_TMA
    // staging queue state: slots are written and read round-robin; a slot's barrier phase flips every time it goes round
    const uint32_t bar0 = (uint32_t) __cvta_generic_to_shared(&sBar[warp][0]);
    int wrIdx = 0, rdIdx = 0;
    uint32_t rdPhase = 0;
    if (STAGED) {
        if (lane == 0) {
#pragma unroll
            for (int k = 0; k < kSlots; ++k) mbar_init(bar0 + 8 * k, 1);
            mbar_fence_init();
        }
        __syncwarp();
    }
#endif

    // descriptor pipeline: tile t + 1's descriptors travel in two registers per lane while tile t is gathered
    int4 nx0 = make_int4(0, 0, 0, 0), nx1 = nx0;
    if (STAGED) {
        const int4* q = reinterpret_cast<const int4*>(P.descs + v0);
        if (lane < V * 4) nx0 = __ldg(q + lane);
        if (lane + 32 < V * 4) nx1 = __ldg(q + lane + 32);
        if (lane < V * 2) reinterpret_cast<int4*>(sVg[warp])[lane] = __ldg(reinterpret_cast<const int4*>(P.vgs + v0) + lane);
    }

    int preV = -1;                                            // ZL_TMA: voice of THIS tile whose window the previous tile already staged
    for (int t = 0; t < P.tilesPerBlock; ++t) {
        const int t0 = t * kT;
        const int n = min(kT, nframes - t0);
        const TileDesc* __restrict__ dp = P.descs + (size_t) t * P.nVoices + v0;
        if (STAGED) {
            __syncwarp();                                     // the previous tile's descriptors are no longer read
            if (lane < V * 4) reinterpret_cast<int4*>(sDesc[warp])[lane] = nx0;
            if (lane + 32 < V * 4) reinterpret_cast<int4*>(sDesc[warp])[lane + 32] = nx1;
            __syncwarp();
            if (t + 1 < P.tilesPerBlock) {
                const int4* q = reinterpret_cast<const int4*>(dp + P.nVoices);
                if (lane < V * 4) nx0 = __ldg(q + lane);
                if (lane + 32 < V * 4) nx1 = __ldg(q + lane + 32);
            }
        }
        float2 acc[kNC];
#pragma unroll
        for (int ch = 0; ch < kNC; ++ch) acc[ch] = make_float2(0.0f, 0.0f);
        // block peak of a voice -> positions model gain  :213-216, :265
        auto take_peak = [&](float vmax, int vs, int flags) {
            if (!JMODE) {
                // The reduction is NOT put under the kWantPeak test: ptxas 12.9 was seen to predicate the CREDUX of that form
                // on an unrelated live predicate (the hoisted passthrough test) while its consumers used the right one, so
                // buses with a non-trivial passthrough took a stale uniform register for the peak.  Unconditional, the
                // instruction carries no predicate at all; the test only guards the bookkeeping.
                const int pk = __reduce_max_sync(FULL, __float_as_int(vmax));   // vmax >= +0: int order == float order
                if (flags & kWantPeak) {
                    if (STAGED || V <= 32) { if (lane == vs) vpk = max(vpk, pk); }
                    else if (lane == 0 && pk > 0)
                        atomicMax(reinterpret_cast<int*>(&P.voices[v0 + vs].posGain), __float_as_int(__fmul_rn(__int_as_float(pk), 0.5f)));
                }
            }
        };
        // STAGED: the tile's window-staged voice-tiles, in slot order, form a queue through the kSlots slots: kSlots - 1
        // windows are in flight behind the one being computed.  act = voices that sound in this tile, toStage = of those,
        // not yet staged.  (cp.async engine: exactly one group is committed per queue position - an empty one when nothing
        // is left to stage - so "all but the newest kSlots - 1 groups have landed" is always the window about to be read.)
        unsigned act = 0, toStage = 0;
        int vs = -1;
#if !ZL_TMA
        uint32_t wrSlot = slot0;
        int rdOff = 0;
#endif
#if ZL_TMA && ZL_PRESTAGE
        int nextV = -1;                                       // voice of the NEXT tile staged from this one (-1: not yet / none)
        bool mayPre = t + 1 < P.tilesPerBlock;
#endif
        auto stage_next = [&]() {
            if (toStage) {
                const TileDesc& sd = sDesc[warp][__ffs(toStage) - 1];
#if ZL_BOUNDS_CHECK
                {
                    const VoiceGather& sg = sVg[warp][__ffs(toStage) - 1];
                    const uintptr_t w0 = (uintptr_t) __double_as_longlong(sd.p0b), w1 = w0 + (uintptr_t) (uint32_t) __double2loint(sd.Rb);
                    const uintptr_t c0 = reinterpret_cast<uintptr_t>(sg.data), c1 = c0 + (uintptr_t) (sg.sampleDuration + 9) * 8u;   // length + 8 stereo frames
                    ZL_CHECK(9, w0 >= c0 && w1 <= c1 && (sg.nch & 0xff) == 2);
                }
#endif
                toStage &= toStage - 1;
#if ZL_TMA
                stage_window_tma(sd, lane, slot0 + wrIdx * kSlotBytes, bar0 + 8 * wrIdx);
                wrIdx = wrIdx == kSlots - 1 ? 0 : wrIdx + 1;
#else
                stage_window(sd, lane16, wrSlot);
                wrSlot = wrSlot == slot0 + (kSlots - 1) * kSlotBytes ? slot0 : wrSlot + kSlotBytes;
#endif
            }
#if ZL_TMA && ZL_PRESTAGE
            else if (mayPre) {
                // Nothing of this tile is left to stage and a slot is free: the queue runs on into the NEXT tile, whose
                // descriptors already travel in nx0 (lane 4 v + q holds quad q of voice v, v < 8) - its first window-staged
                // voice starts its way here, one voice-tile and an epilogue ahead, instead of being waited for at the tile start.
                mayPre = false;
                const bool ring = (lane & 3) == 3 && lane < V * 4 && (nx0.w & kCountMask) != 0 && (nx0.w & kRingBit) != 0;
                const unsigned m = __ballot_sync(FULL, ring);
                if (m) {
                    const int src = __ffs(m) - 3;             // lane 4 v + 1: quad 1 = (p0b, Rb) = window address, (bytes | base)
                    const uint32_t alo = (uint32_t) __shfl_sync(FULL, nx0.x, src), ahi = (uint32_t) __shfl_sync(FULL, nx0.y, src);
                    const uint32_t bytes = (uint32_t) __shfl_sync(FULL, nx0.z, src);
                    stage_window_tma_raw(((uint64_t) ahi << 32) | alo, bytes, slot0 + wrIdx * kSlotBytes, bar0 + 8 * wrIdx);
                    wrIdx = wrIdx == kSlots - 1 ? 0 : wrIdx + 1;
                    nextV = src >> 2;
                }
            }
#endif
#if !ZL_TMA
            cp_async_commit();
#endif
        };
        if (STAGED) {
            const int fl = lane < V ? sDesc[warp][lane].nrend : 0;
            act = __ballot_sync(FULL, (fl & kCountMask) != 0);
            toStage = __ballot_sync(FULL, (fl & kCountMask) != 0 && (fl & kRingBit) != 0);
            bool prime = true;
#if ZL_TMA && ZL_PRESTAGE
            if (preV >= 0) { toStage &= ~(1u << preV); prime = false; }   // already on its way (or landed): the queue is one ahead
#endif
            if (prime) {
#pragma unroll
                for (int k = 0; k < kSlots - 1; ++k) stage_next();
            }
#if ZL_L2PF == 1
            // the windows that will be staged later in this tile start their way from DRAM to L2 now (one bulk prefetch each)
            for (unsigned m = toStage; m; m &= m - 1) {
                const TileDesc& sd = sDesc[warp][__ffs(m) - 1];
                const uint64_t src = (uint64_t) __double_as_longlong(sd.p0b);
                const uint32_t bytes = (uint32_t) __double2loint(sd.Rb);
                asm volatile("{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\n@p cp.async.bulk.prefetch.L2.global [%0], %1;\n}" :: "l"(src), "r"(bytes) : "memory");
            }
#elif ZL_L2PF == 2
            // same, line by line: four lanes per voice, lane part p takes the 128-byte lines p, p + 4, ... of its voice's window
            for (int vb = 0; vb < V; vb += 8) {
                const int pv = vb + (lane >> 2);
                if (pv < V && ((toStage >> pv) & 1u)) {
                    const int4 q = reinterpret_cast<const int4*>(&sDesc[warp][pv])[1];       // p0b = window address, Rb = (base, bytes)
                    const int o = (lane & 3) * 128;
                    const char* a = reinterpret_cast<const char*>(((uint64_t) (uint32_t) q.y << 32) | (uint32_t) q.x) + o;
#pragma unroll
                    for (int k = 0; k < kSlotBytes / 512; ++k)
                        if (o + k * 512 < q.z) asm volatile("prefetch.global.L2 [%0];" :: "l"(a + k * 512) : "memory");
                }
            }
#endif
        }
        for (;;) {
            if (STAGED) { if (!act) break; vs = __ffs(act) - 1; act &= act - 1; }   // silent voices are never visited
            else if (++vs >= V) break;
            const TileDesc d = STAGED ? sDesc[warp][vs] : load_desc(dp + vs);
            const int nr = d.nrend & kCountMask;              // warp-uniform
            if (!STAGED && nr == 0) continue;
            const VoiceGather g = STAGED ? sVg[warp][vs] : load_vg(P.vgs + v0 + vs);
            if (STAGED && (d.nrend & kRingBit)) {
                stage_next();                                 // keep the queue full: the slot read last time is free
#if ZL_TMA
                mbar_wait(bar0 + 8 * rdIdx, rdPhase);
                const unsigned char* win = sWin[warp][0] + rdIdx * kSlotBytes;
                if (rdIdx == kSlots - 1) { rdIdx = 0; rdPhase ^= 1u; } else ++rdIdx;
#else
                cp_async_wait<kSlots - 1>();
                __syncwarp();                                 // the window was copied by all lanes
                const unsigned char* win = sWin[warp][0] + rdOff;
                rdOff = rdOff == (kSlots - 1) * kSlotBytes ? 0 : rdOff + kSlotBytes;
#endif
                float vmax = 0.0f;
#pragma unroll
                for (int gi = 0; gi < kNG; ++gi)
                    vmax = fmaxf(vmax, gather_finish_ring<JMODE, Q1>(d, g, win, laneD, laneF, nz, pm1, gi * kG * 32, acc + gi * kG));
                take_peak(vmax, vs, g.nch);
                __syncwarp();                                 // every lane is done with the slot before it is staged again
                continue;
            }
            float vmax = 0.0f;
#pragma unroll
            for (int gi = 0; gi < kNG; ++gi) {
                const int c0 = gi * kG * 32;
                if (c0 >= nr) break;                          // warp-uniform: nothing of this voice in the rest of the tile
                Taps tp;
                if ((d.nrend & kSimpleBit) && nr >= c0 + kG * 32) {
                    gather_issue_simple<true>(d, g, lane, laneD, laneF, c0, tp);
                    vmax = fmaxf(vmax, gather_finish<JMODE, Q1, true>(g, tp, kG, nz, pm1, acc + gi * kG));
                } else {
                    const int nchunks = min(kG, (min(n, nr) - c0 + 31) >> 5);
                    if (d.nrend & kSimpleBit) gather_issue_simple<false>(d, g, lane, laneD, laneF, c0, tp);
                    else {
                        const size_t vt = (size_t) t * P.nVoices + v0 + vs;
                        gather_issue<JMODE>(d, g, P.te + vt * kT, P.ps + vt * kT, lane, nchunks, c0, tp);
                    }
                    vmax = fmaxf(vmax, gather_finish<JMODE, Q1, false>(g, tp, nchunks, nz, pm1, acc + gi * kG));
                }
            }
            take_peak(vmax, vs, g.nch);
        }
#if ZL_TMA && ZL_PRESTAGE
        preV = nextV;
#endif
        // raw SamplerChannel ports, then JackPassthrough and the meters on the register accumulators
#pragma unroll
        for (int ch = 0; ch < kNC; ++ch) {
            const int c = ch * 32 + lane;
            const int j = t0 + c + QS;                            // output index in the block
            if (LEAN) {
                if (c < n && j < nframes) {
                    float dL, dR;
                    ZL_CHECK(8, t0 + c + QS >= 0 && t0 + c + QS < nframes && off + nframes <= P.stride);
                    if (out0L) { out0L[t0 + c] = acc[ch].x; out0R[t0 + c] = acc[ch].y; }
                    passthrough_pair(pt, pt.dry_amount, acc[ch].x, acc[ch].y, dL, dR);
                    pkL = max(pkL, peak_int(dL)); pkR = max(pkR, peak_int(dR));
                }
            } else if (c < n && j < nframes) {
                float dL, dR;
                ZL_CHECK(8, j >= 0 && j < nframes && off + nframes <= P.stride);
                if (P.out[0]) { P.out[0][rowL + j] = acc[ch].x; P.out[0][rowR + j] = acc[ch].y; }
                passthrough_pair(pt, pt.dry_amount, acc[ch].x, acc[ch].y, dL, dR);
                if (P.out[1]) { P.out[1][rowL + j] = dL; P.out[1][rowR + j] = dR; }
                if (P.out[2]) {
                    float xL, xR;
                    passthrough_pair(pt, pt.wet_fx1_amount, acc[ch].x, acc[ch].y, xL, xR);
                    P.out[2][rowL + j] = xL; P.out[2][rowR + j] = xR;
                }
                if (P.out[3]) {
                    float xL, xR;
                    passthrough_pair(pt, pt.wet_fx2_amount, acc[ch].x, acc[ch].y, xL, xR);
                    P.out[3][rowL + j] = xL; P.out[3][rowR + j] = xR;
                }
                pkL = max(pkL, peak_int(dL)); pkR = max(pkR, peak_int(dR));
                if (P.sumsq) {
                    sqL += (double) __fmul_rn(dL, dL);                 // juce_AudioSampleBuffer.h:1087-1095
                    sqR += (double) __fmul_rn(dR, dR);
                }
            }
        }
    }
    if (!JMODE && (STAGED || V <= 32) && lane < V) {               // the positions-model gain of the voices that take it this period
        const int flags = STAGED ? sVg[warp][lane].nch : __ldg(&P.vgs[v0 + lane].nch);
        if (flags & kWantPeak) P.voices[v0 + lane].posGain = __fmul_rn(__int_as_float(vpk), 0.5f);
    }
    if (Q2 && lane == 0) {                                         // index 0 is never accumulated into (:218): stays zero
#pragma unroll 1
        for (int w = 0; w < 4; ++w)
            if (P.out[w]) { P.out[w][rowL] = 0.0f; P.out[w][rowR] = 0.0f; }
    }
    const size_t idx = ((size_t) blk * P.nSessions * P.B + gbw) * 2;
    if (P.peaks) {
        pkL = __reduce_max_sync(FULL, pkL);
        pkR = __reduce_max_sync(FULL, pkR);
        if (lane == 0) { P.peaks[idx] = pkL; P.peaks[idx + 1] = pkR; }
    }
    if (P.sumsq) {                                                 // fixed order: per-lane partial sums over the block, then the lane tree
#pragma unroll 1
        for (int o = 16; o > 0; o >>= 1) {
            sqL += __shfl_down_sync(FULL, sqL, o);
            sqR += __shfl_down_sync(FULL, sqR, o);
        }
        if (lane == 0) { P.sumsq[idx] = sqL; P.sumsq[idx + 1] = sqR; }
    }
    ZL_TRACE_END(0)
}

// ---------------------------------------------------------------------------------------------
// JACK port mix of each session: master = dry(bus 0), then += dry(bus b) in bus order, where dry is
// JackPassthroughPrivate::process applied to the raw SamplerChannel port (lib/JackPassthrough.cpp:94-101) — recomputed
// here from the raw plane with the same arithmetic the gather used, so the dry plane need not exist.
// One thread per (session, frame); both channels; the raw rows are read fully coalesced.
template <int VEC>
__global__ void __launch_bounds__(256)
master_mix_kernel(const float* __restrict__ raw, const BusDev* __restrict__ buses, float* __restrict__ master,
                  int nSessions, int B, int64_t stride, int64_t n) {
    const int64_t j = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) * VEC;   // VEC consecutive frames per thread (float4 rows)
    const int s = blockIdx.y;
    if (j >= n) return;
    float mL[VEC], mR[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) { mL[k] = 0.0f; mR[k] = 0.0f; }
    for (int b = 0; b < B; ++b) {
        const int gb = s * B + b;
        const zl_passthrough_params pt = buses[gb].pt;
        float inL[VEC], inR[VEC];
        if (VEC == 4) {
            const float4 l = __ldcs(reinterpret_cast<const float4*>(raw + ((int64_t) gb * 2) * stride + j));       // streamed once
            const float4 r = __ldcs(reinterpret_cast<const float4*>(raw + ((int64_t) gb * 2 + 1) * stride + j));
            inL[0] = l.x; inL[1 % VEC] = l.y; inL[2 % VEC] = l.z; inL[3 % VEC] = l.w;
            inR[0] = r.x; inR[1 % VEC] = r.y; inR[2 % VEC] = r.z; inR[3 % VEC] = r.w;
        } else {
            inL[0] = raw[((int64_t) gb * 2) * stride + j]; inR[0] = raw[((int64_t) gb * 2 + 1) * stride + j];
        }
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
            float dL, dR;
            passthrough_pair(pt, pt.dry_amount, inL[k], inR[k], dL, dR);
            if (b == 0) { mL[k] = dL; mR[k] = dR; } else { mL[k] = __fadd_rn(mL[k], dL); mR[k] = __fadd_rn(mR[k], dR); }
        }
    }
    float* oL = master + ((int64_t) s * 2) * stride + j;
    float* oR = oL + stride;
    if (VEC == 4) {
        *reinterpret_cast<float4*>(oL) = make_float4(mL[0], mL[1 % VEC], mL[2 % VEC], mL[3 % VEC]);
        *reinterpret_cast<float4*>(oR) = make_float4(mR[0], mR[1 % VEC], mR[2 % VEC], mR[3 % VEC]);
    } else { oL[0] = mL[0]; oR[0] = mR[0]; }
}

cudaError_t launch_master_mix(const float* raw, const BusDev* buses, float* master, int nSessions, int B, int64_t stride,
                              int64_t n, cudaStream_t stream) {
    if ((n & 3) == 0 && (stride & 3) == 0) {
        const dim3 grid((unsigned) ((n / 4 + 255) / 256), (unsigned) nSessions);
        master_mix_kernel<4><<<grid, 256, 0, stream>>>(raw, buses, master, nSessions, B, stride, n);
    } else {
        const dim3 grid((unsigned) ((n + 255) / 256), (unsigned) nSessions);
        master_mix_kernel<1><<<grid, 256, 0, stream>>>(raw, buses, master, nSessions, B, stride, n);
    }
    return cudaGetLastError();
}

// zl_clip_destroy: voices still on the clip end as by stopNote(0, false) (lib/SamplerSynthVoice.cpp:152-168), so no
// record keeps an id whose PCM is about to be freed.
__global__ void __launch_bounds__(256) stop_clip_voices_kernel(VoiceDev* voices, int n, int clipId) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    VoiceDev& v = voices[i];
    if (v.sound == clipId || v.clip == clipId) voice_stop(v, false);
}
cudaError_t launch_stop_clip_voices(VoiceDev* voices, int n, int clipId, cudaStream_t stream) {
    stop_clip_voices_kernel<<<(n + 255) / 256, 256, 0, stream>>>(voices, n, clipId);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Scratch of one block: descs | vgs | te rows | ps rows (offsets through `off`, 256-byte aligned).
size_t render_scratch_bytes(int nVoices, int nframes, size_t off[5]) {
    const size_t tiles = (size_t) (nframes + kT - 1) / kT;
    auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
    off[0] = 0;
    off[1] = up(off[0] + tiles * nVoices * sizeof(TileDesc));
    off[2] = up(off[1] + (size_t) nVoices * sizeof(VoiceGather));
    off[3] = up(off[2] + tiles * nVoices * kT * sizeof(float));
    off[4] = up(off[3] + tiles * nVoices * kT * sizeof(double));
    return up(off[4] + (size_t) nVoices * sizeof(int32_t));
}

void render_bind_scratch(RenderParams& P, void* scratch, int nVoices, int nframes) {
    size_t off[5];
    render_scratch_bytes(nVoices, nframes, off);
    unsigned char* b = static_cast<unsigned char*>(scratch);
    P.descs = reinterpret_cast<TileDesc*>(b + off[0]);
    P.vgs = reinterpret_cast<VoiceGather*>(b + off[1]);
    P.te = reinterpret_cast<float*>(b + off[2]);
    P.ps = reinterpret_cast<double*>(b + off[3]);
    P.walkQueue = reinterpret_cast<int32_t*>(b + off[4]);
    P.nVoices = nVoices;
    P.tilesPerBlock = (nframes + kT - 1) / kT;
}

// Profiling aid (ZL_B200_DEBUG_SKIP bits 8 / 16, results become wrong): stand-ins for walk_kernel that keep its shape - 625
// one-warp CTAs busy for about as long - but (8) run a dependent FMA chain in a 4-instruction loop, i.e. no instruction-cache
// footprint and no memory traffic, or (16) additionally wander through a large body of straight-line code.
__global__ void __launch_bounds__(32, 1) spin_small_kernel(float* sink, int iters) {
    float x = (float) threadIdx.x, y = 1.0001f;
#pragma unroll 1
    for (int i = 0; i < iters; ++i) x = __fmaf_rn(x, y, 0.5f);
    if (x == 12345.678f) *sink = x;
}
__device__ __forceinline__ float big_body(float x, float y) {
#pragma unroll
    for (int k = 1; k <= 600; ++k) {
        x = __fmaf_rn(x, y, (float) k); x = __fmaf_rn(x, x, 0.25f); x = __fmul_rn(x, 0.5f); x = __fadd_rn(x, (float) (k * 3));
    }
    return x;
}
__global__ void __launch_bounds__(32, 1) spin_big_kernel(float* sink, int iters) {
    float x = (float) threadIdx.x, y = 1.0001f;
#pragma unroll 1
    for (int i = 0; i < iters; ++i) x = big_body(x, y);       // 2400 dependent instructions = 38 KB of code per iteration
    if (x == 12345.678f) *sink = x;
}

// (32): the small spin plus walk-like memory traffic - every lane loads a 224-byte voice record piece by piece (32 sectors per
// instruction) and stores 16-byte pieces of its own 2 KB envelope row, as plan_voice_block / plan_envelope_walk do.
__global__ void __launch_bounds__(32, 1) spin_mem_kernel(const RenderParams P, int iters) {
    const unsigned v = ((blockIdx.x * 32u + threadIdx.x) * 7919u) % (unsigned) P.nVoices;
    const int4* rec = reinterpret_cast<const int4*>(P.voices + v);
    float4* row = reinterpret_cast<float4*>(P.te + (size_t) v * kT);
    float x = (float) threadIdx.x;
    int acc = 0;
#pragma unroll 1
    for (int k = 0; k < 14; ++k) acc += __ldg(rec + k).x;               // the record: 14 x 16 B
#pragma unroll 1
    for (int i = 0; i < iters; ++i) {
#pragma unroll 1
        for (int k = 0; k < 64; ++k) x = __fmaf_rn(x, 1.0001f, 0.5f);
        row[i & 127] = make_float4(x, x, x, (float) acc);               // 8 frames of envelope per 2 stores in the real thing
    }
    if (x == 12345.678f) P.walkQueue[0] = acc;
}

// One block (P.blk) of every session: launch_plan (plan_kernel + walk_kernel), then launch_gather.  The two may sit on
// different streams (zl_api.cu orders them with events and double-buffers the hand-over), so that the latency-bound
// planner of period b + 1 fills the issue slots the gather of period b leaves idle.
#if ZL_BOUNDS_CHECK
void render_dump_trace(const char* path) {
    unsigned long long h[16] = {0};
    cudaDeviceSynchronize();
    if (cudaMemcpyFromSymbol(h, g_bounds, sizeof h) != cudaSuccess) return;
    FILE* f = fopen(path, "a");
    if (!f) return;
    fprintf(f, "bounds checks %llu; violations: table row %llu, tap (generic) %llu, tap (simple) %llu, tap (staged window) %llu, window size/alignment %llu, "
               "queue write %llu, queue length %llu, output index %llu, window vs clip %llu\n", h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7], h[8], h[9]);
    fclose(f);
}
#elif ZL_CTA_TRACE
void render_dump_trace(const char* path) {
    static unsigned long long h[3][kTraceMax][3];
    cudaDeviceSynchronize();
    if (cudaMemcpyFromSymbol(h, g_trace, sizeof h) != cudaSuccess) return;
    FILE* f = fopen(path, "w");
    if (!f) return;
    for (int k = 0; k < 3; ++k)
        for (int i = 0; i < kTraceMax; ++i)
            if (h[k][i][2]) fprintf(f, "%d %d %llu %llu %llu\n", k, i, h[k][i][0], h[k][i][1], h[k][i][2]);
    fclose(f);
}
#else
void render_dump_trace(const char*) {}
#endif

cudaError_t launch_plan(RenderParams& P, cudaStream_t stream) {
    const int V = P.V, nBuses = P.nSessions * P.B;
    if (V < 1 || V > 1024) return cudaErrorInvalidValue;
    static const int planThreadsWanted = [] { const char* e = getenv("ZL_B200_PLAN_THREADS"); const int v = e ? atoi(e) : 64; return v == 32 || v == 128 ? v : 64; }();
    P.planBusesPerCta = V >= planThreadsWanted ? 1 : planThreadsWanted / V;
    const int planThreads = ((P.planBusesPerCta * V + 31) / 32) * 32;
    P.planGroups = (nBuses + P.planBusesPerCta - 1) / P.planBusesPerCta;
    const int planGrid = P.planCtaCap > 0 ? min(P.planGroups, P.planCtaCap) : P.planGroups;
    const bool q3 = (P.quirks & ZL_Q3_RENOTEOFF) != 0;
    // Timing decomposition aid (results become WRONG, never set outside a profiling run): after the first 96 periods the
    // planner kernels named by the mask are no longer launched and the gather replays stale hand-over records.
    static const int dbgSkip = [] { const char* e = getenv("ZL_B200_DEBUG_SKIP"); return e ? atoi(e) : 0; }();
    static int dbgCalls = 0;
    const int skip = (dbgSkip && ++dbgCalls > 96) ? dbgSkip : 0;
    static const int walkProf = [] { const char* e = getenv("ZL_B200_WALK_PROF"); return e && atoi(e) ? 8 : 0; }();
    P.dbgFlags = (skip & 4) | walkProf;
    if (skip & 1) {
    } else if (planThreads <= 128) {
        if (P.mode == ZL_MODE_J) ZL_LAUNCH((plan_kernel<true, false, 128>), planGrid, planThreads, stream, P);
        else if (q3) ZL_LAUNCH((plan_kernel<false, true, 128>), planGrid, planThreads, stream, P);
        else ZL_LAUNCH((plan_kernel<false, false, 128>), planGrid, planThreads, stream, P);
    } else {
        if (P.mode == ZL_MODE_J) ZL_LAUNCH((plan_kernel<true, false, 1024>), planGrid, planThreads, stream, P);
        else if (q3) ZL_LAUNCH((plan_kernel<false, true, 1024>), planGrid, planThreads, stream, P);
        else ZL_LAUNCH((plan_kernel<false, false, 1024>), planGrid, planThreads, stream, P);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    const int wGrid = min((P.nVoices + 31) / 32 + 1, P.walkCtaCap > 0 ? P.walkCtaCap : 148 * 16);
    if (skip & 2) {
        if (skip & 8) spin_small_kernel<<<625, 32, 0, stream>>>(reinterpret_cast<float*>(P.walkQueue), 60000);     // ~60000 x 4 cycles = 120 us
        if (skip & 16) spin_big_kernel<<<625, 32, 0, stream>>>(reinterpret_cast<float*>(P.walkQueue), 25);          // 25 x 2400 x 4 cycles = 120 us
        if (skip & 32) spin_mem_kernel<<<625, 32, 0, stream>>>(P, 128);                                              // 128 x 64 FMAs + 128 scattered 16-byte stores per lane
        return cudaGetLastError();
    }
    if (P.mode == ZL_MODE_J) ZL_LAUNCH((walk_kernel<true, false>), wGrid, 32, stream, P);
    else if (q3) ZL_LAUNCH((walk_kernel<false, true>), wGrid, 32, stream, P);
    else ZL_LAUNCH((walk_kernel<false, false>), wGrid, 32, stream, P);
    return cudaGetLastError();
}

cudaError_t launch_gather(RenderParams& P, cudaStream_t stream, cudaEvent_t gatherBegin, cudaEvent_t gatherEnd) {
    const int V = P.V, nBuses = P.nSessions * P.B;
    cudaError_t e;
    if (gatherBegin && (e = cudaEventRecord(gatherBegin, stream)) != cudaSuccess) return e;
    const int gGrid = (nBuses + ZL_GATHER_THREADS / 32 - 1) / (ZL_GATHER_THREADS / 32);
    const bool q1 = (P.quirks & ZL_Q1_TAP_GAIN) != 0, q2 = (P.quirks & ZL_Q2_PREINC) != 0;
    const bool lean = !P.out[1] && !P.out[2] && !P.out[3] && !P.sumsq;
#define ZL_LAUNCH_GATHER(J, A, B2)                                                                          \
    do {                                                                                                   \
        if (V <= kStageV) {                                                                                \
            if (lean) ZL_LAUNCH((gather_kernel<J, A, B2, true, true>), gGrid, ZL_GATHER_THREADS, stream, P);  \
            else ZL_LAUNCH((gather_kernel<J, A, B2, true, false>), gGrid, ZL_GATHER_THREADS, stream, P);      \
        } else ZL_LAUNCH((gather_kernel<J, A, B2, false, false>), gGrid, ZL_GATHER_THREADS, stream, P);       \
    } while (0)
    if (P.mode == ZL_MODE_J) ZL_LAUNCH_GATHER(true, false, false);
    else if (q1 && q2) ZL_LAUNCH_GATHER(false, true, true);
    else if (q1) ZL_LAUNCH_GATHER(false, true, false);
    else if (q2) ZL_LAUNCH_GATHER(false, false, true);
    else ZL_LAUNCH_GATHER(false, false, false);
#undef ZL_LAUNCH_GATHER
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if (gatherEnd) return cudaEventRecord(gatherEnd, stream);
    return cudaSuccess;
}

}  // namespace zl
