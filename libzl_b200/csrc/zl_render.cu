// zl_render.cu — the batched sampler render kernel (sm_100a).
//
// One CTA renders one session (an independent JACK graph: B SamplerChannels x V voices) for all blocks
// of the launch:  SamplerChannel::process (lib/SamplerSynth.cpp:116-148) -> SamplerSynthVoice::process
// (lib/SamplerSynthVoice.cpp:174-270) or juce::SamplerVoice::renderNextBlock (juce_Sampler.cpp:117-163)
// -> JackPassthroughPrivate::process (lib/JackPassthrough.cpp:45-115) -> AudioLevels peak scan
// (lib/AudioLevels.cpp:357-385) -> JACK port mix of the dry outputs (master).
//
// Phases per (block, bus):  warp 0 drains the bus's commands and walks the V voice state machines for
// the block (lane = voice), leaving per-frame (pos, alpha, env) tables in shared memory; all threads then
// gather + interpolate + pan the voices in slot order (the reference's serial += order) with the bus
// accumulators in registers, and run the passthrough / meter / master epilogue on the same registers.
#include "zl_voice.cuh"

namespace zl {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;

struct __align__(16) RenderSmem {
    float peakPart[kWarps][2];     // per-warp int peaks (as int bits) for the bus epilogue
    double sqPart[kWarps][2];
    int vpeak[32];                 // per-voice block peak (float bits, >= 0)
    VoiceBlock vb[32];
    unsigned long long vframes, vsrc;
};

__device__ __forceinline__ float lerp_gain_z(float a, float b, float ia, float alpha, float g, float env, float vol, bool q1) {
    if (q1)   // lib/SamplerSynthVoice.cpp:204-205 as written: gain/env/volume multiply only the second tap
        return __fadd_rn(__fmul_rn(a, ia), __fmul_rn(__fmul_rn(__fmul_rn(__fmul_rn(b, alpha), g), env), vol));
    return __fmul_rn(__fmul_rn(__fmul_rn(__fadd_rn(__fmul_rn(a, ia), __fmul_rn(b, alpha)), g), env), vol);
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

__global__ void __launch_bounds__(kThreads)
render_kernel(const RenderParams P) {
    extern __shared__ __align__(16) unsigned char smemRaw[];
    const int s = blockIdx.x;
    const int B = P.B, V = P.V, NV = B * V;
    const int nframes = P.nframes;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool jmode = P.mode == ZL_MODE_J;
    const int quirks = jmode ? 0 : P.quirks;
    const bool q1 = (quirks & ZL_Q1_TAP_GAIN) != 0, q2 = (quirks & ZL_Q2_PREINC) != 0;

    // shared memory carve-up
    RenderSmem* sm = reinterpret_cast<RenderSmem*>(smemRaw);
    VoiceDev* sv = reinterpret_cast<VoiceDev*>(smemRaw + ((sizeof(RenderSmem) + 15) & ~size_t(15)));
    float* mst = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(sv) + ((sizeof(VoiceDev) * NV + 15) & ~size_t(15)));
    int* tpos = reinterpret_cast<int*>(mst + 2 * nframes);           // [V][nframes]
    float* talpha = reinterpret_cast<float*>(tpos + V * nframes);
    float* tenv = talpha + V * nframes;

    // stage the session's voice records
    {
        const uint32_t* src = reinterpret_cast<const uint32_t*>(P.voices + (size_t) s * NV);
        uint32_t* dst = reinterpret_cast<uint32_t*>(sv);
        for (int i = tid; i < (int) (sizeof(VoiceDev) * NV / 4); i += kThreads) dst[i] = src[i];
        if (tid == 0) { sm->vframes = 0; sm->vsrc = 0; }
    }
    __syncthreads();

    for (int blk = 0; blk < P.nblocks; ++blk) {
        const zl_block_clock ck = P.clocks[blk];
        const int64_t off = (int64_t) blk * nframes;
        for (int bi = 0; bi < B; ++bi) {
            const int gb = s * B + bi;
            BusDev& bus = P.buses[gb];
            VoiceDev* bv = sv + bi * V;
            // ---- phase W: commands + voice walk (warp 0) -----------------------------------------
            if (warp == 0) {
                if (lane == 0) {                                              // drain the ring  lib/SamplerSynth.cpp:118-122
                    int cur = bus.cmdBegin;
                    const int end = bus.cmdBegin + bus.cmdCount;
                    while (cur < end && P.cmds[cur].block == blk) {
                        bus_handle_command(bv, V, bi - 2, P.cmds[cur], P);
                        ++cur;
                    }
                    bus.cmdCount = end - cur;
                    bus.cmdBegin = cur;
                }
                __syncwarp();
                if (lane < V) {
                    VoiceDev v = bv[lane];                                    // registers for the walk
                    VoiceBlock vb;
                    vb.nrend = 0; vb.srcFloats = 0; vb.data = nullptr; vb.nch = 0; vb.sampleDuration = 0;
                    vb.lgain = vb.rgain = vb.vol = vb.lPan = vb.rPan = 0.0f;
                    int n = 0;
                    if (bus.enabled && (v.flags & VF_PLAYING)) {              // :136-140
                        if (jmode) n = voice_walk_j(v, P, nframes, tpos + lane * nframes, talpha + lane * nframes, tenv + lane * nframes, vb);
                        else n = voice_walk_z(v, P, ck, nframes, tpos + lane * nframes, talpha + lane * nframes, tenv + lane * nframes, vb);
                    }
                    bv[lane] = v;
                    sm->vb[lane] = vb;
                    sm->vpeak[lane] = 0;
                    if (n) { atomicAdd(&sm->vframes, (unsigned long long) n); atomicAdd(&sm->vsrc, (unsigned long long) vb.srcFloats); }
                }
            }
            __syncthreads();
            // ---- phase G: gather + mix + epilogue ------------------------------------------------
            const zl_passthrough_params pt = bus.pt;
            int pkL = 0, pkR = 0;
            double sqL = 0.0, sqR = 0.0;
            // output index j in [0, nframes] ; frame = j - 1 when Q2 (frame f lands at index f+1, :218-221)
            const int jEnd = q2 ? nframes + 1 : nframes;
            for (int j0 = 0; j0 < jEnd; j0 += kThreads) {
                const int j = j0 + tid;
                const int f = q2 ? j - 1 : j;
                float accL = 0.0f, accR = 0.0f;
                for (int vi = 0; vi < V; ++vi) {
                    const VoiceBlock& vb = sm->vb[vi];
                    const int nrend = vb.nrend;
                    if (nrend == 0) continue;                                  // warp-uniform
                    float vpk = 0.0f;
                    if (j < jEnd && f >= 0 && f < nrend) {
                        const int pos = tpos[vi * nframes + f];
                        const float alpha = talpha[vi * nframes + f];
                        const float env = tenv[vi * nframes + f];
                        const float ia = __fsub_rn(1.0f, alpha);
                        float l, r;
                        if (jmode) {                                           // juce_Sampler.cpp:128-152
                            float a0, a1, b0, b1;
                            if (vb.nch > 1) {
                                const float2 a = __ldg(reinterpret_cast<const float2*>(vb.data) + pos);
                                const float2 b = __ldg(reinterpret_cast<const float2*>(vb.data) + pos + 1);
                                a0 = a.x; a1 = a.y; b0 = b.x; b1 = b.y;
                            } else {
                                a0 = a1 = __ldg(vb.data + pos); b0 = b1 = __ldg(vb.data + pos + 1);
                            }
                            l = __fadd_rn(__fmul_rn(a0, ia), __fmul_rn(b0, alpha));
                            r = vb.nch > 1 ? __fadd_rn(__fmul_rn(a1, ia), __fmul_rn(b1, alpha)) : l;
                            l = __fmul_rn(l, __fmul_rn(vb.lgain, env));
                            r = __fmul_rn(r, __fmul_rn(vb.rgain, env));
                        } else {                                               // lib/SamplerSynthVoice.cpp:198-216
                            const bool inRange = vb.sampleDuration > pos;
                            l = 0.0f;
                            if (vb.nch > 1) {
                                float2 a = make_float2(0.f, 0.f), b = a;
                                if (inRange) {
                                    a = __ldg(reinterpret_cast<const float2*>(vb.data) + pos);
                                    b = __ldg(reinterpret_cast<const float2*>(vb.data) + pos + 1);
                                    l = lerp_gain_z(a.x, b.x, ia, alpha, vb.lgain, env, vb.vol, q1);
                                    r = lerp_gain_z(a.y, b.y, ia, alpha, vb.rgain, env, vb.vol, q1);
                                } else r = l;
                            } else {
                                if (inRange) {
                                    const float a = __ldg(vb.data + pos), b = __ldg(vb.data + pos + 1);
                                    l = lerp_gain_z(a, b, ia, alpha, vb.lgain, env, vb.vol, q1);
                                }
                                r = l;
                            }
                            const float m = __fmul_rn(0.5f, __fadd_rn(l, r));   // :208 (0.5*(l+r) in f64 is exact)
                            const float sd = __fsub_rn(l, r);
                            l = __fadd_rn(__fmul_rn(vb.lPan, m), sd);
                            r = __fsub_rn(__fmul_rn(vb.rPan, m), sd);
                            vpk = __fadd_rn(l, r);
                        }
                        accL = __fadd_rn(accL, l);
                        accR = __fadd_rn(accR, r);
                    }
                    if (!jmode) {                                              // block peak of the voice  :213-216
                        int pk = max(0, __float_as_int(vpk));
                        pk = __reduce_max_sync(0xffffffffu, pk);
                        if (lane == 0 && pk > 0) atomicMax(&sm->vpeak[vi], pk);
                    }
                }
                if (j < nframes) {
                    // raw SamplerChannel ports, then JackPassthrough, meters, master
                    if (P.out[0]) {
                        P.out[0][((int64_t) gb * 2) * P.stride + off + j] = accL;
                        P.out[0][((int64_t) gb * 2 + 1) * P.stride + off + j] = accR;
                    }
                    float dL, dR, xL, xR;
                    passthrough_pair(pt, pt.dry_amount, accL, accR, dL, dR);
                    if (P.out[1]) {
                        P.out[1][((int64_t) gb * 2) * P.stride + off + j] = dL;
                        P.out[1][((int64_t) gb * 2 + 1) * P.stride + off + j] = dR;
                    }
                    if (P.out[2]) {
                        passthrough_pair(pt, pt.wet_fx1_amount, accL, accR, xL, xR);
                        P.out[2][((int64_t) gb * 2) * P.stride + off + j] = xL;
                        P.out[2][((int64_t) gb * 2 + 1) * P.stride + off + j] = xR;
                    }
                    if (P.out[3]) {
                        passthrough_pair(pt, pt.wet_fx2_amount, accL, accR, xL, xR);
                        P.out[3][((int64_t) gb * 2) * P.stride + off + j] = xL;
                        P.out[3][((int64_t) gb * 2 + 1) * P.stride + off + j] = xR;
                    }
                    pkL = max(pkL, peak_int(dL));
                    pkR = max(pkR, peak_int(dR));
                    sqL += (double) __fmul_rn(dL, dL);                         // juce_AudioSampleBuffer.h:1087-1095
                    sqR += (double) __fmul_rn(dR, dR);
                    if (bi == 0) { mst[j] = dL; mst[nframes + j] = dR; }       // JACK port mix: copy, then add
                    else { mst[j] = __fadd_rn(mst[j], dL); mst[nframes + j] = __fadd_rn(mst[nframes + j], dR); }
                }
            }
            // bus meters: fixed-order reduction (lane tree, then warps in order)
            if (P.peaks || P.sumsq) {
                pkL = __reduce_max_sync(0xffffffffu, pkL);
                pkR = __reduce_max_sync(0xffffffffu, pkR);
                for (int o = 16; o > 0; o >>= 1) {
                    sqL += __shfl_down_sync(0xffffffffu, sqL, o);
                    sqR += __shfl_down_sync(0xffffffffu, sqR, o);
                }
                if (lane == 0) {
                    sm->peakPart[warp][0] = __int_as_float(pkL); sm->peakPart[warp][1] = __int_as_float(pkR);
                    sm->sqPart[warp][0] = sqL; sm->sqPart[warp][1] = sqR;
                }
            }
            __syncthreads();
            if (tid == 0 && (P.peaks || P.sumsq)) {
                int a = 0, b = 0; double x = 0.0, y = 0.0;
                for (int w = 0; w < kWarps; ++w) {
                    a = max(a, __float_as_int(sm->peakPart[w][0])); b = max(b, __float_as_int(sm->peakPart[w][1]));
                    x += sm->sqPart[w][0]; y += sm->sqPart[w][1];
                }
                const size_t idx = ((size_t) blk * P.nSessions * B + gb) * 2;
                if (P.peaks) { P.peaks[idx] = a; P.peaks[idx + 1] = b; }
                if (P.sumsq) { P.sumsq[idx] = x; P.sumsq[idx + 1] = y; }
            }
            // positions model update   lib/SamplerSynthVoice.cpp:265-267
            if (warp == 0 && lane < V) {
                VoiceDev& v = bv[lane];
                if (sm->vb[lane].data != nullptr && v.clip >= 0 && (jmode || (v.flags & VF_HAS_POSID))) {
                    v.posGain = jmode ? 0.0f : __fmul_rn(__int_as_float(sm->vpeak[lane]), 0.5f);
                    v.posProgress = __double2float_rn(__ddiv_rn(v.p, v.srcLen));
                }
            }
            __syncthreads();
        }
        if (P.master) {
            for (int j = tid; j < nframes; j += kThreads) {
                P.master[((int64_t) s * 2) * P.stride + off + j] = mst[j];
                P.master[((int64_t) s * 2 + 1) * P.stride + off + j] = mst[nframes + j];
            }
        }
        __syncthreads();
    }

    // write the voice records back
    {
        uint32_t* dst = reinterpret_cast<uint32_t*>(P.voices + (size_t) s * NV);
        const uint32_t* src = reinterpret_cast<const uint32_t*>(sv);
        for (int i = tid; i < (int) (sizeof(VoiceDev) * NV / 4); i += kThreads) dst[i] = src[i];
        if (tid == 0 && sm->vframes) {
            atomicAdd(P.voiceFrames, sm->vframes); atomicAdd(P.voiceFrames + 1, sm->vsrc);
            atomicAdd(P.voiceFrames + 2, sm->vframes); atomicAdd(P.voiceFrames + 3, sm->vsrc);
        }
    }
}

size_t render_smem_bytes(int B, int V, int nframes) {
    size_t n = (sizeof(RenderSmem) + 15) & ~size_t(15);
    n += (sizeof(VoiceDev) * (size_t) B * V + 15) & ~size_t(15);
    n += sizeof(float) * 2 * (size_t) nframes;
    n += (size_t) V * nframes * 12;
    return n;
}

cudaError_t launch_render(const RenderParams& P, cudaStream_t stream) {
    const size_t smem = render_smem_bytes(P.B, P.V, P.nframes);
    static size_t configured = 0;
    if (smem > configured) {
        cudaError_t e = cudaFuncSetAttribute(render_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
        if (e != cudaSuccess) return e;
        configured = smem;
    }
    render_kernel<<<P.nSessions, kThreads, smem, stream>>>(P);
    return cudaGetLastError();
}

}  // namespace zl
