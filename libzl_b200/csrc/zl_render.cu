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
// Only the per-session master mix crosses warps (one shared stage + two barriers per tile).
#include "zl_voice.cuh"

namespace zl {

constexpr int kT = 64;               // frames per tile in the regular mode (multiple of 32)
constexpr int kNC = kT / 32;         // 32-frame chunks per tile
constexpr int kShort = 32;           // frames per tile when some lane of the warp needs the per-frame position walk
constexpr int kTblFloats = (32 * (kT + 1) > 32 * 33 * 3) ? 32 * (kT + 1) : 32 * 33 * 3;

// What G needs to know about one voice for one tile (shared memory, 48 bytes = 3 x LDS.128).
struct __align__(16) TileDesc {
    double p0a, Ra, p0b, Rb;         // position segments a: frames [0, cb), b: frames [cb, nrend)
    int32_t cb;
    int32_t nrend;                   // frames of the tile this voice renders
    float envConst;
    int32_t kind;                    // bit 0: envelope from the table row; bit 1: positions from the slot row
};

struct __align__(16) WarpSmem {
    // regular tiles: te[32][kT + 1] envelope rows.  short tiles: te[32][33] followed by ps[32][33] f64 position slots.
    float tbl[kTblFloats];
    TileDesc td[32];
    VoiceGather vg[32];
    int   vpeak[32];                 // per-voice block peak of l + r (float bits, >= 0)   :213-216
    int   busPk[64];                 // per-bus block peaks [bus][2]
    double busSq[64];                // per-bus block sums of squares [bus][2]
};

template <bool Q1>
__device__ __forceinline__ float lerp_gain_z(float a, float b, float ia, float alpha, float g, float env, float vol) {
    if (Q1)   // lib/SamplerSynthVoice.cpp:204-205 as written: gain/env/volume multiply only the second tap
        return __fadd_rn(__fmul_rn(a, ia), __fmul_rn(__fmul_rn(__fmul_rn(__fmul_rn(b, alpha), g), env), vol));
    return __fmul_rn(__fmul_rn(__fmul_rn(__fadd_rn(__fmul_rn(a, ia), __fmul_rn(b, alpha)), g), env), vol);
}

// One voice-frame: both taps, interpolation, gains, (Z) M/S pan.  J: juce_Sampler.cpp:128-152, Z: :198-211
template <bool JMODE, bool Q1, bool STEREO>
__device__ __forceinline__ void gather_voice(const VoiceGather& g, int pos, float alpha, float env, float& l, float& r) {
    const float ia = __fsub_rn(1.0f, alpha);
    if (JMODE) {
        if (STEREO) {
            const float2 a = __ldg(reinterpret_cast<const float2*>(g.data) + pos);
            const float2 b = __ldg(reinterpret_cast<const float2*>(g.data) + pos + 1);
            l = __fadd_rn(__fmul_rn(a.x, ia), __fmul_rn(b.x, alpha));
            r = __fadd_rn(__fmul_rn(a.y, ia), __fmul_rn(b.y, alpha));
        } else {
            const float a = __ldg(g.data + pos), b = __ldg(g.data + pos + 1);
            l = __fadd_rn(__fmul_rn(a, ia), __fmul_rn(b, alpha));
            r = l;
        }
        const float ge = __fmul_rn(g.gain, env);
        l = __fmul_rn(l, ge);
        r = __fmul_rn(r, ge);
        return;
    }
    l = 0.0f; r = 0.0f;
    if (g.sampleDuration > pos) {
        if (STEREO) {
            const float2 a = __ldg(reinterpret_cast<const float2*>(g.data) + pos);
            const float2 b = __ldg(reinterpret_cast<const float2*>(g.data) + pos + 1);
            l = lerp_gain_z<Q1>(a.x, b.x, ia, alpha, g.gain, env, g.vol);
            r = lerp_gain_z<Q1>(a.y, b.y, ia, alpha, g.gain, env, g.vol);
        } else {
            const float a = __ldg(g.data + pos), b = __ldg(g.data + pos + 1);
            l = lerp_gain_z<Q1>(a, b, ia, alpha, g.gain, env, g.vol);
            r = l;
        }
    }
    const float m = __fmul_rn(0.5f, __fadd_rn(l, r));   // :208 (0.5*(l+r) in f64 is exact)
    const float sd = __fsub_rn(l, r);
    l = __fadd_rn(__fmul_rn(g.lPan, m), sd);
    r = __fsub_rn(__fmul_rn(g.rPan, m), sd);
}

// One voice x one tile in G: lane = frame within each 32-frame chunk.  Everything the branches test is warp-uniform.
template <bool JMODE, bool Q1, bool STEREO>
__device__ __forceinline__ float gather_tile(const TileDesc& d, const VoiceGather& g, const float* terow, const double* psrow,
                                             int lane, double laneD, float (&accL)[kNC], float (&accR)[kNC]) {
    float vmax = 0.0f;
    const bool single = d.cb >= d.nrend && !(d.kind & 2);
#pragma unroll
    for (int ch = 0; ch < kNC; ++ch) {
        const int c = ch * 32 + lane;
        if (c < d.nrend) {
            double p;
            if (single) p = __fma_rn(__dadd_rn(laneD, (double) (ch * 32)), d.Ra, d.p0a);
            else if (d.kind & 2) p = psrow[c];
            else {
                const bool sb = c >= d.cb;
                p = __fma_rn((double) (c - (sb ? d.cb : 0)), sb ? d.Rb : d.Ra, sb ? d.p0b : d.p0a);
            }
            // (int) p and p - (int) p without int<->f64 conversions: p + 2^52 rounded down is floor(p) + 2^52
            const double t = __dadd_rd(p, 4503599627370496.0);
            const int pos = __double2loint(t);
            const float alpha = __double2float_rn(__dsub_rn(p, __dsub_rn(t, 4503599627370496.0)));
            const float env = (d.kind & 1) ? terow[c] : d.envConst;
            float l, r;
            gather_voice<JMODE, Q1, STEREO>(g, pos, alpha, env, l, r);
            accL[ch] = __fadd_rn(accL[ch], l);
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
// MAXT: 256 (sessions of up to 256 voices, <= 128 registers) or 1024.
template <int CFG, int MAXT>
__global__ void __launch_bounds__(MAXT, MAXT == 256 ? 2 : 1)
render_kernel(const RenderParams P) {
    constexpr bool JMODE = (CFG >> 3) == ZL_MODE_J;
    constexpr bool Q1 = (CFG & ZL_Q1_TAP_GAIN) != 0, Q2 = (CFG & ZL_Q2_PREINC) != 0, Q3 = (CFG & ZL_Q3_RENOTEOFF) != 0;
    constexpr int QS = Q2 ? 1 : 0;                        // frame f lands at output index f + QS   :218-221
    extern __shared__ __align__(16) unsigned char smemRaw[];
    const int s = blockIdx.x;
    const int B = P.B, V = P.V, NV = B * V;
    const int nframes = P.nframes;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nwarps = blockDim.x >> 5;
    const unsigned FULL = 0xffffffffu;

    WarpSmem& ws = reinterpret_cast<WarpSmem*>(smemRaw)[warp];
    float* stage = reinterpret_cast<float*>(smemRaw + sizeof(WarpSmem) * nwarps);   // [B][2][kT] master staging

    // ownership: this lane's voice / bus; the warp's buses
    const bool hasVoice = tid < NV;
    const int bi = tid / V, slot = tid - bi * V;          // bus within the session, voice slot within the bus
    const int gb = s * B + bi;
    const int busesPerWarp = 32 / V;
    const int warpBus0 = warp * busesPerWarp;
    VoiceDev* rec = P.voices + (size_t) s * NV + tid;

    VoiceHot h;
    h.p = 0.0; h.ratio = 0.0; h.relDen = 1.0; h.srcLen = 1.0; h.nextLoopTick = 0; h.nextLoopUsecs = 0;
    h.env = 0.f; h.rRate = 0.f; h.aRate = 0.f; h.dRate = 0.f; h.sustain = 0.f; h.release = 0.f; h.gain = 0.f;
    h.posGain = 0.f; h.posProgress = 0.f; h.adsrState = kIdle; h.flags = 0; h.sound = -1; h.clip = -1; h.slice = -1; h.looping = 0;
    int cmdCur = 0, cmdEnd = 0, busEnabled = 0;
    if (hasVoice) {
        hot_load(h, *rec);
        const BusDev& bus = P.buses[gb];
        cmdCur = bus.cmdBegin; cmdEnd = cmdCur + bus.cmdCount; busEnabled = bus.enabled;
    }
    unsigned long long nVoiceFrames = 0, nSrcFloats = 0;
    const double laneD = (double) lane;

    for (int blk = 0; blk < P.nblocks; ++blk) {
        const zl_block_clock ck = P.clocks[blk];
        const int64_t off = (int64_t) blk * nframes;

        // ---- commands of this block: SamplerChannel::process ring drain   lib/SamplerSynth.cpp:118-122 ----------
        const bool hasCmd = hasVoice && cmdCur < cmdEnd && P.cmds[cmdCur].block == blk;
        if (__any_sync(FULL, hasCmd)) {
            if (hasCmd) hot_store(h, *rec);
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
            if (hasCmd) hot_load(h, *rec);
        }

        // ---- block prelude -------------------------------------------------------------------------------
        VoiceBlockCtl bc;
        bc.stopD = 0.0; bc.startD = 0.0; bc.relStart = 0.0; bc.usPerFrame = 0.0; bc.lengthInTicks = 0; bc.fWrap = nframes;
        bc.beatAligned = 0; bc.spanStart = 0; bc.lastPos = 0; bc.span = 0; bc.nch = 0; bc.clipFrames = 0; bc.releaseArmed = 0;
        bool active = false;
        {
            VoiceGather g;
            g.data = nullptr; g.sampleDuration = 0; g.nch = 0; g.gain = g.vol = g.lPan = g.rPan = 0.0f;
            if (hasVoice && busEnabled && (h.flags & VF_PLAYING))                       // :136-140
                active = voice_block_begin<JMODE>(h, P, ck, nframes, bc, g);
            ws.vg[lane] = g;
            ws.vpeak[lane] = 0;
            ws.busPk[lane] = 0; ws.busPk[lane + 32] = 0;
            ws.busSq[lane] = 0.0; ws.busSq[lane + 32] = 0.0;
        }
        const bool walked = active;
        int nrendBlock = 0;
        __syncwarp();

        for (int t0 = 0; t0 < nframes; t0 += kT) {
            // ---- W: every lane plans its voice's tile ----------------------------------------------------
            const int nT = min(kT, nframes - t0);
            PosPlan pl;
            pl.p0a = 0.0; pl.Ra = 0.0; pl.p0b = 0.0; pl.Rb = 0.0; pl.cb = nT; pl.nPos = 0; pl.kRel = INT_MAX; pl.stopped = 0; pl.irregular = 0;
            PlanUndo undo;
            plan_save(undo, h, bc);
            if (active) plan_positions<JMODE>(h, bc, ck, t0, nT, nframes, pl);
            // a lane that cannot express its tile in closed form makes the warp cover it in 32-frame sub-tiles
            const bool shortTile = __any_sync(FULL, active && pl.irregular);
            if (shortTile && active) plan_restore(undo, h, bc);
            const int nn = shortTile ? kShort : nT;
            const int teStride = shortTile ? 33 : kT + 1;
            float* te = ws.tbl + lane * teStride;
            double* ps = reinterpret_cast<double*>(ws.tbl + 32 * 33) + lane * 33;
          for (int s0 = 0; s0 < nT; s0 += nn) {
            const int n = min(nn, nT - s0);
            if (shortTile && active) {
                plan_save(undo, h, bc);
                plan_positions<JMODE>(h, bc, ck, t0 + s0, n, nframes, pl);
                if (pl.irregular) {
                    plan_restore(undo, h, bc);
                    walk_positions<JMODE>(h, bc, ck, t0 + s0, n, nframes, ps, pl);
                }
            }
            int cnt = 0;
            {
                TileDesc d;
                d.p0a = pl.p0a; d.Ra = pl.Ra; d.p0b = pl.p0b; d.Rb = pl.Rb; d.cb = pl.cb; d.nrend = 0; d.envConst = 0.0f; d.kind = 0;
                if (active) {
                    bool fromTable, ended;
                    cnt = plan_envelope<JMODE, Q3>(h, bc, pl, te, d.envConst, fromTable, ended);
                    d.nrend = cnt;
                    d.kind = (fromTable ? 1 : 0) | (pl.irregular ? 2 : 0);
                    if (ended && cnt < pl.nPos) {                          // the ADSR ran out before the tile did: p is frame cnt's
                        if (pl.irregular) h.p = ps[cnt];
                        else h.p = cnt >= pl.cb ? __fma_rn((double) (cnt - pl.cb), pl.Rb, pl.p0b) : __fma_rn((double) cnt, pl.Ra, pl.p0a);
                    }
                    if (pl.stopped || ended) { voice_stop(h, false); active = false; }
                    else prefetch_span_l2(ws.vg[lane].data, bc.nch, bc.clipFrames, h.p, h.ratio, kT);
                }
                ws.td[lane] = d;
            }
            nrendBlock += cnt;
            __syncwarp();

            // ---- G: lane = frame -----------------------------------------------------------------------------
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
                    if (d.nrend <= 0) continue;                                // warp-uniform: voice silent in this tile
                    const VoiceGather g = ws.vg[row];
                    const float* terow = ws.tbl + row * teStride;
                    const double* psrow = reinterpret_cast<const double*>(ws.tbl + 32 * 33) + row * 33;
                    const float vmax = g.nch > 1 ? gather_tile<JMODE, Q1, true>(d, g, terow, psrow, lane, laneD, accL, accR)
                                                 : gather_tile<JMODE, Q1, false>(d, g, terow, psrow, lane, laneD, accL, accR);
                    if (!JMODE) {                                              // block peak of the voice  :213-216
                        const int pk = __reduce_max_sync(FULL, __float_as_int(vmax));   // vmax >= +0: int order == float order
                        if (lane == 0) atomicMax(&ws.vpeak[row], pk);
                    }
                }
                // raw SamplerChannel ports, then JackPassthrough, meters, master staging
                const zl_passthrough_params pt = P.buses[gbw].pt;
                int pkL = 0, pkR = 0;
                double sqL = 0.0, sqR = 0.0;
#pragma unroll
                for (int ch = 0; ch < kNC; ++ch) {
                    const int c = ch * 32 + lane;
                    const int j = t0 + s0 + c + QS;                            // output index in the block
                    float dL = 0.0f, dR = 0.0f;
                    if (c < n && j < nframes) {
                        const int64_t o = off + j;
                        if (P.out[0]) {
                            P.out[0][((int64_t) gbw * 2) * P.stride + o] = accL[ch];
                            P.out[0][((int64_t) gbw * 2 + 1) * P.stride + o] = accR[ch];
                        }
                        passthrough_pair(pt, pt.dry_amount, accL[ch], accR[ch], dL, dR);
                        if (P.out[1]) {
                            P.out[1][((int64_t) gbw * 2) * P.stride + o] = dL;
                            P.out[1][((int64_t) gbw * 2 + 1) * P.stride + o] = dR;
                        }
                        float xL, xR;
                        if (P.out[2]) {
                            passthrough_pair(pt, pt.wet_fx1_amount, accL[ch], accR[ch], xL, xR);
                            P.out[2][((int64_t) gbw * 2) * P.stride + o] = xL;
                            P.out[2][((int64_t) gbw * 2 + 1) * P.stride + o] = xR;
                        }
                        if (P.out[3]) {
                            passthrough_pair(pt, pt.wet_fx2_amount, accL[ch], accR[ch], xL, xR);
                            P.out[3][((int64_t) gbw * 2) * P.stride + o] = xL;
                            P.out[3][((int64_t) gbw * 2 + 1) * P.stride + o] = xR;
                        }
                        pkL = max(pkL, peak_int(dL)); pkR = max(pkR, peak_int(dR));
                        sqL += (double) __fmul_rn(dL, dL);                     // juce_AudioSampleBuffer.h:1087-1095
                        sqR += (double) __fmul_rn(dR, dR);
                    }
                    if (P.master && c < nn) { stage[((size_t) wb * 2) * kT + s0 + c] = dL; stage[((size_t) wb * 2 + 1) * kT + s0 + c] = dR; }
                }
                if (Q2 && t0 + s0 == 0 && lane == 0) {                              // index 0 is never accumulated into (:218): stays zero
                    for (int w = 0; w < 4; ++w)
                        if (P.out[w]) { P.out[w][((int64_t) gbw * 2) * P.stride + off] = 0.0f; P.out[w][((int64_t) gbw * 2 + 1) * P.stride + off] = 0.0f; }
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
            // ---- JACK port mix of the session (sum of the dry outs in bus order) ------------------------------
            if (P.master) {
                __syncthreads();
                for (int t2 = tid; t2 < 2 * kT; t2 += blockDim.x) {
                    const int ch = t2 / kT, c = t2 - ch * kT;
                    const float* st = stage + ch * kT + c;
                    float m = st[0];
                    for (int b2 = 1; b2 < B; ++b2) m = __fadd_rn(m, st[(size_t) b2 * 2 * kT]);
                    const int j = t0 + c + QS;
                    if (c < nT && j < nframes) P.master[((int64_t) s * 2 + ch) * P.stride + off + j] = m;
                    if (Q2 && t0 == 0 && c == 0) P.master[((int64_t) s * 2 + ch) * P.stride + off] = 0.0f;
                }
                __syncthreads();
            }
        }

        // ---- block epilogue ----------------------------------------------------------------------------------
        if (walked) {
            nVoiceFrames += (unsigned long long) nrendBlock;
            if (nrendBlock > 0) nSrcFloats += (unsigned long long) (bc.span + (uint32_t) (bc.lastPos + 2 - bc.spanStart)) * (unsigned) bc.nch;
            // positions model update   lib/SamplerSynthVoice.cpp:265-267
            if (h.clip >= 0 && (JMODE || (h.flags & VF_HAS_POSID))) {
                h.posGain = JMODE ? 0.0f : __fmul_rn(__int_as_float(ws.vpeak[lane]), 0.5f);
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
    if (hasVoice) {
        hot_store(h, *rec);
        if (slot == 0) { BusDev& bus = P.buses[gb]; bus.cmdBegin = cmdCur; bus.cmdCount = cmdEnd - cmdCur; }
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
static int render_threads(int B, int V) { return ((B * V + 31) / 32) * 32; }

size_t render_smem_bytes(int B, int V, int nframes) {
    (void) nframes;
    const int nv = B * V;
    if (nv > 1024) return (size_t) 1 << 30;                 // one thread per voice of the session: reported as "does not fit"
    return sizeof(WarpSmem) * (size_t) (render_threads(B, V) / 32) + sizeof(float) * (size_t) B * 2 * kT;
}

template <int CFG, int MAXT>
static cudaError_t launch_cfg_t(const RenderParams& P, cudaStream_t stream) {
    const size_t smem = render_smem_bytes(P.B, P.V, P.nframes);
    static size_t configured = 0;
    if (smem > configured) {
        cudaError_t e = cudaFuncSetAttribute(render_kernel<CFG, MAXT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
        if (e != cudaSuccess) return e;
        configured = smem;
    }
    render_kernel<CFG, MAXT><<<P.nSessions, render_threads(P.B, P.V), smem, stream>>>(P);
    return cudaGetLastError();
}

template <int CFG>
static cudaError_t launch_cfg(const RenderParams& P, cudaStream_t stream) {
    return render_threads(P.B, P.V) <= 256 ? launch_cfg_t<CFG, 256>(P, stream) : launch_cfg_t<CFG, 1024>(P, stream);
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
