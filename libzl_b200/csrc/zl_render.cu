// zl_render.cu — the batched sampler render kernel (sm_100a).
//
// One CTA renders one session (an independent JACK graph: B SamplerChannels x V voices) for all blocks of
// the launch:  SamplerChannel::process (lib/SamplerSynth.cpp:116-148) -> SamplerSynthVoice::process
// (lib/SamplerSynthVoice.cpp:174-270) or juce::SamplerVoice::renderNextBlock (juce_Sampler.cpp:117-163)
// -> JackPassthroughPrivate::process (lib/JackPassthrough.cpp:45-115) -> AudioLevels peak scan
// (lib/AudioLevels.cpp:357-385) -> JACK port mix of the dry outputs (master).
//
// Work decomposition (stage 3).  Thread t of the CTA owns voice t of the session for the whole launch
// (bus = t / V, slot = t % V); its state machine lives in registers.  A warp therefore owns 32 voices =
// 32/V whole buses, and everything up to the bus outputs is warp-local:
//   W  each lane PLANS its voice's next tile of kT frames in O(1): the f64 position recurrence p += ratio is
//      exact-linear inside a binade (zl_voice.cuh), so a tile is at most two (p0, R) segments broken at binade
//      crossings / loop wraps / the stop position; the envelope is a constant while the ADSR sits in sustain and
//      is stepped frame by frame into a shared-memory row only while it attacks, decays or releases.  Tiles a
//      voice cannot express that way (p < 1, rounding ties, tiny loops) make the warp drop to 32-frame tiles in
//      which that lane walks its positions frame by frame into a shared-memory slot;
//   G  the warp turns 90 degrees: lane = frame.  For each of its buses it walks the voices in slot order (the
//      serial += order of the reference), evaluates p = p0 + k*R with one DFMA, splits it into (pos, alpha),
//      gathers both taps from the HBM-resident clip (lanes read consecutive source frames -> coalesced), applies
//      gain / envelope / M-S pan, keeps the bus accumulators in registers and runs the passthrough + meter
//      epilogue on them; 128-byte coalesced stores.
// Warps never wait for each other: the per-session master mix (the only cross-bus step) is master_mix_kernel below,
// a streaming pass over the raw bus planes.
#include "zl_voice.cuh"

namespace zl {

#ifndef ZL_MINB
#define ZL_MINB 2
#endif
#ifndef ZL_KT
#define ZL_KT 128
#endif
#ifndef ZL_VPW_SMALL
#define ZL_VPW_SMALL 16
#endif
constexpr int kT = ZL_KT;            // frames per tile in the regular mode (multiple of 32)
constexpr int kNC = kT / 32;         // 32-frame chunks per tile
constexpr int kShort = 32;           // frames per tile when some lane of the warp needs the per-frame position walk

// What G needs to know about one voice for one tile (shared memory, 64 bytes = 4 x LDS.128).
struct __align__(16) TileDesc {
    double p0a, Ra, p0b, Rb;         // position segments a: frames [0, cb), b: frames [cb, nrend)
    float e0a, Ea, e0b, Eb;          // envelope segments a: frames [0, ce), b: frames [ce, ct); frames >= ct: table row
    int32_t cb, ce, ct;
    int32_t nrend;                   // frames of the tile this voice renders; bit 30: positions from the slot row
};
constexpr int kSlotBit = 1 << 30;
constexpr int kSimpleBit = 1 << 29;   // one position segment, one envelope segment, stereo, every tap inside the sample
constexpr int kCountMask = (1 << 29) - 1;

// VPW = voices per warp (16 or 32): fewer voices per warp = more warps per session = more latency hiding.
template <int VPW>
struct __align__(16) WarpSmem {
    // regular tiles: te[VPW][kT + 1] envelope rows.  short tiles: te[VPW][33] followed by ps[VPW][33] f64 position slots.
    static constexpr int kTblFloats = (VPW * (kT + 1) > VPW * 33 * 3) ? VPW * (kT + 1) : VPW * 33 * 3;
    float tbl[kTblFloats];
    TileDesc td[VPW];
    VoiceGather vg[VPW];
    VoiceHot hot[VPW];               // the voices' state machines between tiles (registers only while a lane plans)
    VoiceBlockCtl ctl[VPW];
    unsigned long long cntFrames[VPW], cntSrc[VPW];   // voice-frames / algorithmic source floats of the launch
    int   busPk[64];                 // per-bus block peaks [bus][2]
    double busSq[64];                // per-bus block sums of squares [bus][2]
};

template <bool Q1>
__device__ __forceinline__ float lerp_gain_z(float a, float b, float ia, float alpha, float g, float env, float vol) {
    if (Q1)   // lib/SamplerSynthVoice.cpp:204-205 as written: gain/env/volume multiply only the second tap
        return __fadd_rn(__fmul_rn(a, ia), __fmul_rn(__fmul_rn(__fmul_rn(__fmul_rn(b, alpha), g), env), vol));
    return __fmul_rn(__fmul_rn(__fmul_rn(__fadd_rn(__fmul_rn(a, ia), __fmul_rn(b, alpha)), g), env), vol);
}

// The gather of one voice x one tile is split in two so all of the tile's loads are in flight together:
// gather_issue computes (pos, alpha, env) per lane-frame and issues both tap loads of every chunk; gather_finish does
// the arithmetic.  Lanes past the voice's last frame read zeros (from the clip's zero padding or not at all), so they add
// +0 to the bus and to the voice peak: no validity predicate is needed downstream.
struct Taps {
    float2 a[kNC], b[kNC];           // taps at pos and pos + 1 (mono: .y duplicates .x, so r == l falls out of the same math)
    float alpha[kNC], env[kNC];
};

// (int) p and p - (int) p without int<->f64 conversions: p + 2^52 rounded down is floor(p) + 2^52
__device__ __forceinline__ int split_pos(double p, float& alpha) {
    const double fl = __dadd_rd(p, 4503599627370496.0);
    alpha = __double2float_rn(__dsub_rn(p, __dsub_rn(fl, 4503599627370496.0)));
    return __double2loint(fl);
}

template <bool JMODE>
__device__ __forceinline__ void gather_issue(const TileDesc& d, const VoiceGather& g, const float* terow, const double* psrow,
                                             int lane, int nchunks, Taps& t) {
    const int nrend = d.nrend & kCountMask;
    const bool slotPos = (d.nrend & kSlotBit) != 0;
#pragma unroll
    for (int ch = 0; ch < kNC; ++ch) {
        t.a[ch] = make_float2(0.0f, 0.0f); t.b[ch] = t.a[ch]; t.alpha[ch] = 0.0f; t.env[ch] = 0.0f;
        if (ch < nchunks) {                                           // warp-uniform
            const int c = ch * 32 + lane;
            const bool v = c < nrend;
            double p;
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
                if (g.nch > 1) {
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
__device__ __forceinline__ void gather_issue_simple(const TileDesc& d, const VoiceGather& g, int lane, double laneD, float laneF, Taps& t) {
    const int nrend = d.nrend & kCountMask;
    const float2* src = reinterpret_cast<const float2*>(g.data);
    const int zeroPos = g.sampleDuration + 2;                         // length + 1 (and + 2): inside the padding
    const bool envConst = d.Ea == 0.0f;                               // warp-uniform (sustain, J-mode idle)
#pragma unroll
    for (int ch = 0; ch < kNC; ++ch) {
        const double p = __fma_rn(__dadd_rn(laneD, (double) (ch * 32)), d.Ra, d.p0a);
        int pos = split_pos(p, t.alpha[ch]);
        if (!FULL) pos = (ch * 32 + lane < nrend) ? pos : zeroPos;
        t.a[ch] = __ldg(src + pos); t.b[ch] = __ldg(src + pos + 1);
    }
    if (envConst) {
#pragma unroll
        for (int ch = 0; ch < kNC; ++ch) t.env[ch] = d.e0a;
    } else {
#pragma unroll
        for (int ch = 0; ch < kNC; ++ch) t.env[ch] = __fmaf_rn(__fadd_rn(laneF, (float) (ch * 32)), d.Ea, d.e0a);
    }
}

// Interpolation, gains, (Z) M/S pan, bus accumulate.  J: juce_Sampler.cpp:128-152, Z: lib/SamplerSynthVoice.cpp:198-221
template <bool JMODE, bool Q1>
__device__ __forceinline__ float gather_finish(const VoiceGather& g, const Taps& t, int nchunks, float (&accL)[kNC], float (&accR)[kNC]) {
    float vmax = 0.0f;
#pragma unroll
    for (int ch = 0; ch < kNC; ++ch) {
        if (ch < nchunks) {                                           // warp-uniform
            const float alpha = t.alpha[ch], env = t.env[ch];
            const float ia = __fsub_rn(1.0f, alpha);
            float l, r;
            if (JMODE) {
                l = __fadd_rn(__fmul_rn(t.a[ch].x, ia), __fmul_rn(t.b[ch].x, alpha));
                r = __fadd_rn(__fmul_rn(t.a[ch].y, ia), __fmul_rn(t.b[ch].y, alpha));
                const float ge = __fmul_rn(g.gain, env);
                l = __fmul_rn(l, ge);
                r = __fmul_rn(r, ge);
            } else {
                l = lerp_gain_z<Q1>(t.a[ch].x, t.b[ch].x, ia, alpha, g.gain, env, g.vol);
                r = lerp_gain_z<Q1>(t.a[ch].y, t.b[ch].y, ia, alpha, g.gain, env, g.vol);
                const float m = __fmul_rn(0.5f, __fadd_rn(l, r));   // :208 (0.5*(l+r) in f64 is exact)
                const float sd = __fsub_rn(l, r);
                l = __fadd_rn(__fmul_rn(g.lPan, m), sd);
                r = __fsub_rn(__fmul_rn(g.rPan, m), sd);
            }
            accL[ch] = __fadd_rn(accL[ch], l);                      // lanes past the voice's end add +0
            accR[ch] = __fadd_rn(accR[ch], r);
            vmax = fmaxf(vmax, __fadd_rn(l, r));
        }
    }
    return vmax;
}

// Fire-and-forget L2 prefetch of the source span the voice will read in its NEXT tile (one bulk request per voice:
// UBLKPF).  The gather's loads then hit L2 instead of waiting on HBM; the HBM requests of 32 voices x all resident
// warps are in flight while the current tile is being computed.
__device__ __forceinline__ void prefetch_span_l2(const float* data, int nch, int clipFrames, double p, double ratio, int frames) {
    const int first = (int) p;
    int last = first + (int) (ratio * (double) frames) + 2;
    last = min(last, clipFrames + 7);
    if (last <= first) return;
    const size_t fb = (size_t) nch * sizeof(float);
    uintptr_t a = reinterpret_cast<uintptr_t>(data) + (size_t) first * fb;
    uintptr_t e = reinterpret_cast<uintptr_t>(data) + (size_t) (last + 1) * fb;
    a &= ~uintptr_t(15);
    const unsigned bytes = (unsigned) ((e - a) & ~uintptr_t(15));
    if (bytes == 0) return;
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(a), "r"(bytes) : "memory");
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

// CFG = mode << 3 | quirks (quirks forced to 0 in J-mode): compile-time so the frame loops carry no mode tests.
// MAXT: 256 threads (3 CTAs per SM, <= 85 registers) or 1024.
template <int CFG, int MAXT, int VPW>
__global__ void __launch_bounds__(MAXT, MAXT == 256 ? ZL_MINB : 1)
render_kernel(const RenderParams P) {
    constexpr bool JMODE = (CFG >> 3) == ZL_MODE_J;
    constexpr bool Q1 = (CFG & ZL_Q1_TAP_GAIN) != 0, Q2 = (CFG & ZL_Q2_PREINC) != 0, Q3 = (CFG & ZL_Q3_RENOTEOFF) != 0;
    constexpr int QS = Q2 ? 1 : 0;                        // frame f lands at output index f + QS   :218-221
    extern __shared__ __align__(16) unsigned char smemRaw[];
    const int s = blockIdx.x;
    const int B = P.B, V = P.V, NV = B * V;
    const int nframes = P.nframes;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned FULL = 0xffffffffu;

    WarpSmem<VPW>& ws = reinterpret_cast<WarpSmem<VPW>*>(smemRaw)[warp];
    // ownership: this lane's voice / bus; the warp's buses
    const int vIdx = warp * VPW + lane;                   // lanes >= VPW own no voice; they only help in G
    const bool ownsRow = lane < VPW;
    const bool hasVoice = ownsRow && vIdx < NV;
    const int bi = vIdx / V, slot = vIdx - bi * V;        // bus within the session, voice slot within the bus
    const int gb = s * B + bi;
    const int busesPerWarp = VPW / V;
    const int warpBus0 = warp * busesPerWarp;
    VoiceDev* rec = P.voices + (size_t) s * NV + vIdx;
    const double laneD = (double) lane;
    const float laneF = (float) lane;

    int cmdCur = 0, cmdEnd = 0, busEnabled = 0;
    if (ownsRow) {
        VoiceHot h;
        h.p = 0.0; h.ratio = 0.0; h.relDen = 1.0; h.srcLen = 1.0; h.nextLoopTick = 0; h.nextLoopUsecs = 0;
        h.env = 0.f; h.rRate = 0.f; h.aRate = 0.f; h.dRate = 0.f; h.sustain = 0.f; h.release = 0.f; h.gain = 0.f;
        h.posGain = 0.f; h.posProgress = 0.f; h.adsrState = kIdle; h.flags = 0; h.sound = -1; h.clip = -1; h.slice = -1; h.looping = 0;
        if (hasVoice) {
            hot_load(h, *rec);
            if (P.cmdRange) { cmdCur = P.cmdRange[2 * gb]; cmdEnd = cmdCur + P.cmdRange[2 * gb + 1]; }
            busEnabled = P.buses[gb].enabled;
        }
        ws.hot[lane] = h;
        ws.cntFrames[lane] = 0; ws.cntSrc[lane] = 0;
    }

    for (int blk = 0; blk < P.nblocks; ++blk) {
        const int64_t off = (int64_t) blk * nframes;

        // ---- commands of this block: SamplerChannel::process ring drain   lib/SamplerSynth.cpp:118-122 ----------
        const bool hasCmd = hasVoice && cmdCur < cmdEnd && P.cmds[cmdCur].block == blk;
        if (__any_sync(FULL, hasCmd)) {
            if (hasCmd) hot_store(ws.hot[lane], *rec);
            __syncwarp();
            if (hasCmd) {
                int cur = cmdCur;
                while (cur < cmdEnd && P.cmds[cur].block == blk) {
                    if (slot == 0) bus_handle_command(rec, V, bi - 2, P.cmds[cur], P);   // the bus's V records start at slot 0
                    ++cur;
                }
                cmdCur = cur;
            }
            __syncwarp();
            if (hasCmd) hot_load(ws.hot[lane], *rec);
        }

        // ---- block prelude -------------------------------------------------------------------------------
        bool active = false;
        if (ownsRow) {
            const zl_block_clock ck = P.clocks[blk];
            VoiceHot h = ws.hot[lane];
            VoiceBlockCtl bc;
            bc.stopD = 0.0; bc.startD = 0.0; bc.relStart = 0.0; bc.usPerFrame = 0.0; bc.lengthInTicks = 0; bc.fWrap = nframes;
            bc.beatAligned = 0; bc.spanStart = 0; bc.lastPos = 0; bc.span = 0; bc.nch = 0; bc.clipFrames = 0; bc.releaseArmed = 0;
            VoiceGather g;
            g.data = nullptr; g.sampleDuration = 0; g.nch = 0; g.gain = g.vol = g.lPan = g.rPan = 0.0f;
            if (hasVoice && busEnabled && (h.flags & VF_PLAYING))                       // :136-140
                active = voice_block_begin<JMODE>(h, P, ck, nframes, bc, g);
            ws.hot[lane] = h; ws.ctl[lane] = bc;
            ws.vg[lane] = g;
        }
        ws.busPk[lane] = 0; ws.busPk[lane + 32] = 0;
        ws.busSq[lane] = 0.0; ws.busSq[lane + 32] = 0.0;
        const bool walked = active;
        int nrendBlock = 0;
        int vpk = 0;                                          // block peak of this lane's voice, l + r as float bits (>= +0)   :213-216
        __syncwarp();

        for (int t0 = 0; t0 < nframes; t0 += kT) {
            const int nT = min(kT, nframes - t0);
            // ---- W: every owner lane plans its voice's tile (state comes from / returns to shared memory) ---------
            // One O(1) plan per lane per tile.  A lane that cannot express its tile in closed form makes the warp cover
            // the tile in 32-frame sub-tiles instead (every lane re-plans from its saved state).
            VoiceHot h;
            VoiceBlockCtl bc;
            PosPlan pl;
            bool irregular = false;
            if (active) {
                const zl_block_clock ck = P.clocks[blk];
                h = ws.hot[lane]; bc = ws.ctl[lane];
                plan_positions<JMODE>(h, bc, ck, t0, nT, nframes, pl);
                irregular = pl.irregular != 0;
            }
            const bool shortTile = __any_sync(FULL, irregular);
            const int nn = shortTile ? kShort : nT;
            const int teStride = shortTile ? 33 : kT + 1;
          for (int s0 = 0; s0 < nT; s0 += nn) {
            const int n = min(nn, nT - s0);
            const int nchunks = (n + 31) >> 5;
            int cnt = 0;
            if (ownsRow) {
                TileDesc d;
                d.p0a = 0.0; d.Ra = 0.0; d.p0b = 0.0; d.Rb = 0.0; d.e0a = 0.0f; d.Ea = 0.0f; d.e0b = 0.0f; d.Eb = 0.0f;
                d.cb = n; d.ce = n; d.ct = INT_MAX; d.nrend = 0;
                if (active) {
                    float* te = ws.tbl + lane * teStride;
                    double* ps = reinterpret_cast<double*>(ws.tbl + VPW * 33) + lane * 33;
                    if (shortTile) {
                        const zl_block_clock ck = P.clocks[blk];
                        h = ws.hot[lane]; bc = ws.ctl[lane];
                        PlanUndo undo;
                        plan_save(undo, h, bc);
                        plan_positions<JMODE>(h, bc, ck, t0 + s0, n, nframes, pl);
                        if (pl.irregular) {
                            plan_restore(undo, h, bc);
                            walk_positions<JMODE>(h, bc, ck, t0 + s0, n, nframes, ps, pl);
                        }
                    }
                    d.p0a = pl.p0a; d.Ra = pl.Ra; d.p0b = pl.p0b; d.Rb = pl.Rb; d.cb = pl.cb;
                    EnvPlan ep;
                    bool ended;
                    cnt = plan_envelope<JMODE, Q3>(h, bc, pl, te, ep, ended);
                    d.e0a = ep.e0a; d.Ea = ep.Ea; d.e0b = ep.e0b; d.Eb = ep.Eb; d.ce = ep.ce; d.ct = ep.ct;
                    d.nrend = cnt | (pl.irregular ? kSlotBit : 0);
                    if (cnt > 0 && !pl.irregular && pl.cb >= cnt && ep.ce >= cnt && ep.ct >= cnt && bc.nch == 2
                        && (JMODE || (int) __fma_rn((double) (cnt - 1), pl.Ra, pl.p0a) < ws.vg[lane].sampleDuration))
                        d.nrend |= kSimpleBit;
                    if (ended && cnt < pl.nPos) {                          // the ADSR ran out before the tile did: p is frame cnt's
                        if (pl.irregular) h.p = ps[cnt];
                        else h.p = cnt >= pl.cb ? __fma_rn((double) (cnt - pl.cb), pl.Rb, pl.p0b) : __fma_rn((double) cnt, pl.Ra, pl.p0a);
                    }
                    if (pl.stopped || ended) { voice_stop(h, false); active = false; }
                    else prefetch_span_l2(ws.vg[lane].data, bc.nch, bc.clipFrames, h.p, h.ratio, kT);
                    ws.hot[lane] = h; ws.ctl[lane] = bc;
                }
                ws.td[lane] = d;
            }
            nrendBlock += cnt;
            __syncwarp();

            // ---- G: lane = frame; one voice at a time, all chunks of the tile in flight ------------------------------
            const double* psbase = reinterpret_cast<const double*>(ws.tbl + VPW * 33);
            for (int bl = 0; bl < busesPerWarp; ++bl) {
                const int wb = warpBus0 + bl;               // bus of the session
                if (wb >= B) break;
                const int gbw = s * B + wb;
                float accL[kNC], accR[kNC];
#pragma unroll
                for (int ch = 0; ch < kNC; ++ch) { accL[ch] = 0.0f; accR[ch] = 0.0f; }
                for (int vs = 0; vs < V; ++vs) {
                    const int row = bl * V + vs;
                    const TileDesc d = ws.td[row];
                    const int nr = d.nrend & kCountMask;              // warp-uniform
#ifdef ZL_EXP_NOG
                    continue;
#endif
                    if (nr == 0) continue;
                    const VoiceGather g = ws.vg[row];
                    Taps tp;
                    float vmax;
                    if (d.nrend & kSimpleBit) {
                        if (nr == kT) {
                            gather_issue_simple<true>(d, g, lane, laneD, laneF, tp);
                            vmax = gather_finish<JMODE, Q1>(g, tp, kNC, accL, accR);
                        } else {
                            gather_issue_simple<false>(d, g, lane, laneD, laneF, tp);
                            vmax = gather_finish<JMODE, Q1>(g, tp, nchunks, accL, accR);
                        }
                    } else {
                        gather_issue<JMODE>(d, g, ws.tbl + row * teStride, psbase + row * 33, lane, nchunks, tp);
                        vmax = gather_finish<JMODE, Q1>(g, tp, nchunks, accL, accR);
                    }
                    if (!JMODE) {                                              // block peak of the voice  :213-216
                        const int pk = __reduce_max_sync(FULL, __float_as_int(vmax));   // vmax >= +0: int order == float order
                        if (lane == row) vpk = max(vpk, pk);
                    }
                }
                // raw SamplerChannel ports, then JackPassthrough and the meters on the register accumulators
                const zl_passthrough_params pt = P.buses[gbw].pt;
                const int64_t rowL = ((int64_t) gbw * 2) * P.stride + off, rowR = rowL + P.stride;
                int pkL = 0, pkR = 0;
                double sqL = 0.0, sqR = 0.0;
#pragma unroll
                for (int ch = 0; ch < kNC; ++ch) {
                    const int c = ch * 32 + lane;
                    const int j = t0 + s0 + c + QS;                            // output index in the block
                    if (c < n && j < nframes) {
                        float dL, dR;
                        if (P.out[0]) { P.out[0][rowL + j] = accL[ch]; P.out[0][rowR + j] = accR[ch]; }
                        passthrough_pair(pt, pt.dry_amount, accL[ch], accR[ch], dL, dR);
                        if (P.out[1]) { P.out[1][rowL + j] = dL; P.out[1][rowR + j] = dR; }
                        float xL, xR;
                        if (P.out[2]) {
                            passthrough_pair(pt, pt.wet_fx1_amount, accL[ch], accR[ch], xL, xR);
                            P.out[2][rowL + j] = xL; P.out[2][rowR + j] = xR;
                        }
                        if (P.out[3]) {
                            passthrough_pair(pt, pt.wet_fx2_amount, accL[ch], accR[ch], xL, xR);
                            P.out[3][rowL + j] = xL; P.out[3][rowR + j] = xR;
                        }
                        pkL = max(pkL, peak_int(dL)); pkR = max(pkR, peak_int(dR));
                        if (P.sumsq) {
                            sqL += (double) __fmul_rn(dL, dL);                 // juce_AudioSampleBuffer.h:1087-1095
                            sqR += (double) __fmul_rn(dR, dR);
                        }
                    }
                }
                if (Q2 && t0 + s0 == 0 && lane == 0) {                         // index 0 is never accumulated into (:218): stays zero
                    for (int w = 0; w < 4; ++w)
                        if (P.out[w]) { P.out[w][rowL] = 0.0f; P.out[w][rowR] = 0.0f; }
                }
                if (P.peaks) {
                    pkL = __reduce_max_sync(FULL, pkL);
                    pkR = __reduce_max_sync(FULL, pkR);
                    if (lane == 0) { ws.busPk[bl * 2] = max(ws.busPk[bl * 2], pkL); ws.busPk[bl * 2 + 1] = max(ws.busPk[bl * 2 + 1], pkR); }
                }
                if (P.sumsq) {                                                 // fixed order: lane tree, then tiles in order
                    for (int o = 16; o > 0; o >>= 1) {
                        sqL += __shfl_down_sync(FULL, sqL, o);
                        sqR += __shfl_down_sync(FULL, sqR, o);
                    }
                    if (lane == 0) { ws.busSq[bl * 2] += sqL; ws.busSq[bl * 2 + 1] += sqR; }
                }
            }
            __syncwarp();
          }
        }

        // ---- block epilogue ----------------------------------------------------------------------------------
        if (walked) {
            VoiceHot& h = ws.hot[lane];
            const VoiceBlockCtl& bc = ws.ctl[lane];
            ws.cntFrames[lane] += (unsigned long long) nrendBlock;
            if (nrendBlock > 0) ws.cntSrc[lane] += (unsigned long long) (bc.span + (uint32_t) (bc.lastPos + 2 - bc.spanStart)) * (unsigned) bc.nch;
            // positions model update   lib/SamplerSynthVoice.cpp:265-267
            if (h.clip >= 0 && (JMODE || (h.flags & VF_HAS_POSID))) {
                h.posGain = JMODE ? 0.0f : __fmul_rn(__int_as_float(vpk), 0.5f);
                h.posProgress = __double2float_rn(__ddiv_rn(h.p, h.srcLen));
            }
        }
        if (lane < busesPerWarp && warpBus0 + lane < B) {
            const size_t idx = ((size_t) blk * P.nSessions * B + (size_t) s * B + warpBus0 + lane) * 2;
            if (P.peaks) { P.peaks[idx] = ws.busPk[lane * 2]; P.peaks[idx + 1] = ws.busPk[lane * 2 + 1]; }
            if (P.sumsq) { P.sumsq[idx] = ws.busSq[lane * 2]; P.sumsq[idx + 1] = ws.busSq[lane * 2 + 1]; }
        }
        __syncwarp();
    }

    // ---- write the voice records back, account the launch -------------------------------------------------------
    unsigned long long nVoiceFrames = 0, nSrcFloats = 0;
    if (hasVoice) {
        hot_store(ws.hot[lane], *rec);
        nVoiceFrames = ws.cntFrames[lane]; nSrcFloats = ws.cntSrc[lane];
    }
    for (int o = 16; o > 0; o >>= 1) {
        nVoiceFrames += __shfl_down_sync(FULL, nVoiceFrames, o);
        nSrcFloats += __shfl_down_sync(FULL, nSrcFloats, o);
    }
    if (lane == 0 && nVoiceFrames) {
        atomicAdd(P.voiceFrames, nVoiceFrames); atomicAdd(P.voiceFrames + 1, nSrcFloats);
        atomicAdd(P.voiceFrames + 2, nVoiceFrames); atomicAdd(P.voiceFrames + 3, nSrcFloats);
    }
}

// ---------------------------------------------------------------------------------------------
// JACK port mix of each session: master = dry(bus 0), then += dry(bus b) in bus order, where dry is
// JackPassthroughPrivate::process applied to the raw SamplerChannel port (lib/JackPassthrough.cpp:94-101) — recomputed
// here from the raw plane with the same arithmetic the render kernel used, so the dry plane need not exist.
// One thread per (session, frame); both channels; the raw rows are read fully coalesced.
__global__ void __launch_bounds__(256)
master_mix_kernel(const float* __restrict__ raw, const BusDev* __restrict__ buses, float* __restrict__ master,
                  int nSessions, int B, int64_t stride, int64_t n) {
    const int64_t j = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    const int s = blockIdx.y;
    if (j >= n) return;
    float mL = 0.0f, mR = 0.0f;
    for (int b = 0; b < B; ++b) {
        const int gb = s * B + b;
        const zl_passthrough_params pt = buses[gb].pt;
        const float inL = raw[((int64_t) gb * 2) * stride + j], inR = raw[((int64_t) gb * 2 + 1) * stride + j];
        float dL, dR;
        passthrough_pair(pt, pt.dry_amount, inL, inR, dL, dR);
        if (b == 0) { mL = dL; mR = dR; } else { mL = __fadd_rn(mL, dL); mR = __fadd_rn(mR, dR); }
    }
    master[((int64_t) s * 2) * stride + j] = mL;
    master[((int64_t) s * 2 + 1) * stride + j] = mR;
}

cudaError_t launch_master_mix(const float* raw, const BusDev* buses, float* master, int nSessions, int B, int64_t stride,
                              int64_t n, cudaStream_t stream) {
    const dim3 grid((unsigned) ((n + 255) / 256), (unsigned) nSessions);
    master_mix_kernel<<<grid, 256, 0, stream>>>(raw, buses, master, nSessions, B, stride, n);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
static int voices_per_warp(int B, int V) { return (ZL_VPW_SMALL == 16 && V <= 16 && B * V <= 512) ? 16 : 32; }
static int render_threads(int B, int V) { const int w = voices_per_warp(B, V); return ((B * V + w - 1) / w) * 32; }

size_t render_smem_bytes(int B, int V, int nframes) {
    (void) nframes;
    const int nv = B * V;
    if (nv > 1024) return (size_t) 1 << 30;                 // reported as "does not fit"
    const size_t per = voices_per_warp(B, V) == 16 ? sizeof(WarpSmem<16>) : sizeof(WarpSmem<32>);
    return per * (size_t) (render_threads(B, V) / 32);
}

template <int CFG, int MAXT, int VPW>
static cudaError_t launch_cfg_t(const RenderParams& P, cudaStream_t stream) {
    const size_t smem = render_smem_bytes(P.B, P.V, P.nframes);
    static size_t configured = 0;
    if (smem > configured) {
        cudaError_t e = cudaFuncSetAttribute(render_kernel<CFG, MAXT, VPW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
        if (e != cudaSuccess) return e;
        configured = smem;
    }
    render_kernel<CFG, MAXT, VPW><<<P.nSessions, render_threads(P.B, P.V), smem, stream>>>(P);
    return cudaGetLastError();
}

template <int CFG>
static cudaError_t launch_cfg(const RenderParams& P, cudaStream_t stream) {
    const int t = render_threads(P.B, P.V);
    if (voices_per_warp(P.B, P.V) == 16) return t <= 256 ? launch_cfg_t<CFG, 256, 16>(P, stream) : launch_cfg_t<CFG, 1024, 16>(P, stream);
    return t <= 256 ? launch_cfg_t<CFG, 256, 32>(P, stream) : launch_cfg_t<CFG, 1024, 32>(P, stream);
}

cudaError_t launch_render(const RenderParams& P, cudaStream_t stream) {
    if (P.mode == ZL_MODE_J) return launch_cfg<(ZL_MODE_J << 3)>(P, stream);
    switch (P.quirks & 7) {
        case 0: return launch_cfg<0>(P, stream);
        case 1: return launch_cfg<1>(P, stream);
        case 2: return launch_cfg<2>(P, stream);
        case 3: return launch_cfg<3>(P, stream);
        case 4: return launch_cfg<4>(P, stream);
        case 5: return launch_cfg<5>(P, stream);
        case 6: return launch_cfg<6>(P, stream);
        default: return launch_cfg<7>(P, stream);
    }
}

}  // namespace zl
