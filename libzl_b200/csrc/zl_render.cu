// zl_render.cu — the batched sampler render kernel (sm_100a).
//
// One CTA renders one session (an independent JACK graph: B SamplerChannels x V voices) for all blocks of
// the launch:  SamplerChannel::process (lib/SamplerSynth.cpp:116-148) -> SamplerSynthVoice::process
// (lib/SamplerSynthVoice.cpp:174-270) or juce::SamplerVoice::renderNextBlock (juce_Sampler.cpp:117-163)
// -> JackPassthroughPrivate::process (lib/JackPassthrough.cpp:45-115) -> AudioLevels peak scan
// (lib/AudioLevels.cpp:357-385) -> JACK port mix of the dry outputs (master).
//
// Work decomposition (stage 2).  Thread t of the CTA owns voice t of the session for the whole launch
// (bus = t / V, slot = t % V); its state machine lives in registers.  A warp therefore owns 32 voices =
// 32/V whole buses, and everything up to the bus outputs is warp-local:
//   W  each lane steps ITS voice through the next tile of 32 frames with the reference's exact recurrences
//      (f64 position, f32 ADSR, loop / release / stop events) and leaves (pos, alpha, env) per frame in the
//      warp's shared-memory tile table (row = voice, 33 columns: conflict-free both ways);
//   G  the warp turns 90 degrees: lane = frame.  For each of its buses it walks the voices in slot order
//      (the serial += order of the reference), gathers both interpolation taps from the HBM-resident clip
//      (lanes read consecutive source frames -> coalesced), applies gain / envelope / M-S pan, keeps the bus
//      accumulators in registers and runs the passthrough + meter epilogue on them; 128-byte coalesced stores.
// Only the per-session master mix crosses warps (one double-buffered shared stage + one barrier per tile).
#include "zl_voice.cuh"

namespace zl {

constexpr int kTile = 32;            // frames per tile == lanes per warp
constexpr int kCols = kTile + 1;     // column 0 carries the previous tile's last frame (quirk Q2 reads frame j-1)

struct __align__(16) WarpSmem {
    int   tp[32 * kCols];            // (int) sourceSamplePosition per voice-frame
    float ta[32 * kCols];            // alpha
    float te[32 * kCols];            // ADSR envelope value
    VoiceGather vg[32];
    int   rng[32];                   // valid table columns of the voice in this tile: lo | hi << 8
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
template <bool JMODE, bool Q1>
__device__ __forceinline__ void gather_voice(const VoiceGather& g, int pos, float alpha, float env, float& l, float& r) {
    const float ia = __fsub_rn(1.0f, alpha);
    if (JMODE) {
        float a0, a1, b0, b1;
        if (g.nch > 1) {
            const float2 a = __ldg(reinterpret_cast<const float2*>(g.data) + pos);
            const float2 b = __ldg(reinterpret_cast<const float2*>(g.data) + pos + 1);
            a0 = a.x; a1 = a.y; b0 = b.x; b1 = b.y;
        } else {
            a0 = a1 = __ldg(g.data + pos); b0 = b1 = __ldg(g.data + pos + 1);
        }
        l = __fadd_rn(__fmul_rn(a0, ia), __fmul_rn(b0, alpha));
        r = g.nch > 1 ? __fadd_rn(__fmul_rn(a1, ia), __fmul_rn(b1, alpha)) : l;
        const float ge = __fmul_rn(g.gain, env);
        l = __fmul_rn(l, ge);
        r = __fmul_rn(r, ge);
        return;
    }
    const bool inRange = g.sampleDuration > pos;
    l = 0.0f; r = 0.0f;
    if (inRange) {
        if (g.nch > 1) {
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

// CFG = mode << 3 | quirks (quirks forced to 0 in J-mode): compile-time so the frame loops carry no mode tests.
// MAXT: 256 (sessions of up to 256 voices, <= 128 registers) or 1024.
template <int CFG, int MAXT>
__global__ void __launch_bounds__(MAXT, MAXT == 256 ? 2 : 1)
render_kernel(const RenderParams P) {
    constexpr bool JMODE = (CFG >> 3) == ZL_MODE_J;
    constexpr bool Q1 = (CFG & ZL_Q1_TAP_GAIN) != 0, Q2 = (CFG & ZL_Q2_PREINC) != 0, Q3 = (CFG & ZL_Q3_RENOTEOFF) != 0;
    extern __shared__ __align__(16) unsigned char smemRaw[];
    const int s = blockIdx.x;
    const int B = P.B, V = P.V, NV = B * V;
    const int nframes = P.nframes;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nwarps = blockDim.x >> 5;
    const unsigned FULL = 0xffffffffu;

    WarpSmem& ws = reinterpret_cast<WarpSmem*>(smemRaw)[warp];
    float* stage = reinterpret_cast<float*>(smemRaw + sizeof(WarpSmem) * nwarps);   // [2][B][2][32] master staging

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
    int stageBuf = 0;

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
        bc.beatAligned = 0; bc.spanStart = 0; bc.lastPos = 0; bc.span = 0; bc.nch = 0; bc.releaseArmed = 0;
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
        bool carryValid = false;
        int nrendBlock = 0;
        int* tp = ws.tp + lane * kCols;
        float* ta = ws.ta + lane * kCols;
        float* te = ws.te + lane * kCols;
        __syncwarp();

        for (int j0 = 0; j0 < nframes; j0 += kTile) {
            const int n = min(kTile, nframes - j0);
            // ---- W: owner lanes step their voices through the tile --------------------------------------
            int cnt = 0;
            if (active) {
                if (JMODE) cnt = voice_walk_tile_j(h, bc, n, tp + 1, ta + 1, te + 1, active);
                else cnt = voice_walk_tile_z<Q3>(h, bc, ck, j0, n, nframes, tp + 1, ta + 1, te + 1, active);
            }
            nrendBlock += cnt;
            ws.rng[lane] = ((Q2 && carryValid) ? 0 : 1) | ((1 + cnt) << 8);
            __syncwarp();

            // ---- G: lane = frame --------------------------------------------------------------------------
            const int j = j0 + lane;                        // output index in the block
            const int col = Q2 ? lane : lane + 1;           // Q2: frame f lands at index f + 1   :218-221
            for (int bl = 0; bl < busesPerWarp; ++bl) {
                const int wb = warpBus0 + bl;               // bus of the session
                if (wb >= B) break;
                const int gbw = s * B + wb;
                float accL = 0.0f, accR = 0.0f;
                for (int vs = 0; vs < V; ++vs) {
                    const int row = bl * V + vs;
                    const int rg = ws.rng[row];
                    const int lo = rg & 0xff, hi = rg >> 8;
                    if (hi <= lo) continue;                                    // warp-uniform: voice silent in this tile
                    const VoiceGather g = ws.vg[row];
                    float vpk = 0.0f;
                    if (col >= lo && col < hi) {
                        float l, r;
                        gather_voice<JMODE, Q1>(g, ws.tp[row * kCols + col], ws.ta[row * kCols + col], ws.te[row * kCols + col], l, r);
                        accL = __fadd_rn(accL, l);
                        accR = __fadd_rn(accR, r);
                        vpk = __fadd_rn(l, r);
                    }
                    if (!JMODE) {                                              // block peak of the voice  :213-216
                        int pk = max(0, __float_as_int(vpk));
                        pk = __reduce_max_sync(FULL, pk);
                        if (lane == 0 && pk > ws.vpeak[row]) ws.vpeak[row] = pk;
                    }
                }
                // raw SamplerChannel ports, then JackPassthrough, meters, master staging
                const zl_passthrough_params pt = P.buses[gbw].pt;
                float dL = 0.0f, dR = 0.0f;
                int pkL = 0, pkR = 0;
                double sqL = 0.0, sqR = 0.0;
                if (j < nframes) {
                    if (P.out[0]) {
                        P.out[0][((int64_t) gbw * 2) * P.stride + off + j] = accL;
                        P.out[0][((int64_t) gbw * 2 + 1) * P.stride + off + j] = accR;
                    }
                    passthrough_pair(pt, pt.dry_amount, accL, accR, dL, dR);
                    if (P.out[1]) {
                        P.out[1][((int64_t) gbw * 2) * P.stride + off + j] = dL;
                        P.out[1][((int64_t) gbw * 2 + 1) * P.stride + off + j] = dR;
                    }
                    float xL, xR;
                    if (P.out[2]) {
                        passthrough_pair(pt, pt.wet_fx1_amount, accL, accR, xL, xR);
                        P.out[2][((int64_t) gbw * 2) * P.stride + off + j] = xL;
                        P.out[2][((int64_t) gbw * 2 + 1) * P.stride + off + j] = xR;
                    }
                    if (P.out[3]) {
                        passthrough_pair(pt, pt.wet_fx2_amount, accL, accR, xL, xR);
                        P.out[3][((int64_t) gbw * 2) * P.stride + off + j] = xL;
                        P.out[3][((int64_t) gbw * 2 + 1) * P.stride + off + j] = xR;
                    }
                    pkL = peak_int(dL); pkR = peak_int(dR);
                    sqL = (double) __fmul_rn(dL, dL);                          // juce_AudioSampleBuffer.h:1087-1095
                    sqR = (double) __fmul_rn(dR, dR);
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
                if (P.master) {
                    float* st = stage + ((size_t) (stageBuf * B + wb) * 2) * kTile;
                    st[lane] = dL; st[kTile + lane] = dR;
                }
            }
            __syncwarp();
            // ---- carry the tile's last frame into column 0 for the next tile (Q2) ----------------------------
            if (Q2) {
                tp[0] = tp[kTile]; ta[0] = ta[kTile]; te[0] = te[kTile];
                carryValid = cnt == kTile;
            }
            // ---- JACK port mix of the session (sum of the dry outs in bus order) ------------------------------
            if (P.master) {
                __syncthreads();
                for (int t2 = tid; t2 < 2 * kTile; t2 += blockDim.x) {
                    const int ch = t2 >> 5, ln = t2 & 31;
                    const float* st = stage + (size_t) stageBuf * B * 2 * kTile + ch * kTile + ln;
                    float m = st[0];
                    for (int b2 = 1; b2 < B; ++b2) m = __fadd_rn(m, st[(size_t) b2 * 2 * kTile]);
                    if (j0 + ln < nframes) P.master[((int64_t) s * 2 + ch) * P.stride + off + j0 + ln] = m;
                }
                stageBuf ^= 1;
            }
        }

        // ---- block epilogue ----------------------------------------------------------------------------------
        if (!JMODE && Q2 && (nframes % kTile) == 0 && nrendBlock == nframes) {
            // The frame rendered last lands at index nframes (one past the JACK buffer, :218-221): no output lane
            // took it, but it still counts for the voice's block peak (:213-216).  Its table entry is in column 32.
            float l, r;
            gather_voice<JMODE, Q1>(ws.vg[lane], tp[kTile], ta[kTile], te[kTile], l, r);
            const int pk = max(0, __float_as_int(__fadd_rn(l, r)));
            if (pk > ws.vpeak[lane]) ws.vpeak[lane] = pk;
        }
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
    return sizeof(WarpSmem) * (size_t) (render_threads(B, V) / 32) + sizeof(float) * 2 * (size_t) B * 2 * kTile;
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
