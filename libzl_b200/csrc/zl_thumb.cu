// zl_thumb.cu — WaveFormItem / juce::AudioThumbnail reductions and the standalone passthrough / meter
// stages as sm_100a kernels.  All outputs are integers (int8 min/max pairs, int32 peaks, uint8 mask):
// bit-exact against the reference arithmetic.  JUCE/ = tracktion_engine/modules/juce/modules/.
#include "zl_device.cuh"

namespace zl {

// AudioThumbnail::MinMaxValue::setFloat   JUCE/juce_audio_utils/gui/juce_AudioThumbnail.cpp:46-66
__device__ __forceinline__ void set_float(float mn, float mx, int8_t* out) {
    int a = juce_round_to_int((double) __fmul_rn(mn, 127.0f));
    int b = juce_round_to_int((double) __fmul_rn(mx, 127.0f));
    a = max(-128, min(127, a));
    b = max(-128, min(127, b));
    int8_t v0 = (int8_t) a, v1 = (int8_t) b;
    if (v0 == v1) { if (v1 == 127) v0--; else v1++; }
    out[0] = v0; out[1] = v1;
}

// Level 0 (K6): one warp per thumb sample; LevelDataSource::readNextBlock -> readMaxLevels -> setFloat
// (juce_AudioThumbnail.cpp:225-269, JUCE/juce_audio_formats/format/juce_AudioFormatReader.cpp:199-249).
// data is interleaved (float2 frames for stereo).  Only i < nDone = len/spt entries are computed; the
// trailing entry keeps its (0,0) default (sizing :666-670).  out: [nch][nThumb][2].
template <int NCH>
__global__ void __launch_bounds__(256)
thumb_level0_kernel(const float* __restrict__ data, int spt, int nDone, int nThumb, int8_t* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int warpsPerGrid = (gridDim.x * blockDim.x) >> 5;
    for (int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < nDone; i += warpsPerGrid) {
        float mn0 = INFINITY, mx0 = -INFINITY, mn1 = INFINITY, mx1 = -INFINITY;
        const size_t base = (size_t) i * spt * NCH;              // float index of the block start
        const int nfl = spt * NCH;                                // floats in the block
        if ((base & 3) == 0 && (nfl & 3) == 0) {
            const float4* p = reinterpret_cast<const float4*>(data + base);
            const int n4 = nfl >> 2;
            #pragma unroll 4
            for (int k = lane; k < n4; k += 32) {
                const float4 v = __ldcs(p + k);
                if (NCH == 2) {
                    mn0 = fminf(mn0, fminf(v.x, v.z)); mx0 = fmaxf(mx0, fmaxf(v.x, v.z));
                    mn1 = fminf(mn1, fminf(v.y, v.w)); mx1 = fmaxf(mx1, fmaxf(v.y, v.w));
                } else {
                    mn0 = fminf(mn0, fminf(fminf(v.x, v.y), fminf(v.z, v.w)));
                    mx0 = fmaxf(mx0, fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w)));
                }
            }
        } else {
            for (int k = lane; k < spt; k += 32) {
                if (NCH == 2) {
                    const float2 v = __ldcs(reinterpret_cast<const float2*>(data + base) + k);
                    mn0 = fminf(mn0, v.x); mx0 = fmaxf(mx0, v.x);
                    mn1 = fminf(mn1, v.y); mx1 = fmaxf(mx1, v.y);
                } else {
                    const float v = __ldcs(data + base + k);
                    mn0 = fminf(mn0, v); mx0 = fmaxf(mx0, v);
                }
            }
        }
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn0 = fminf(mn0, __shfl_xor_sync(0xffffffffu, mn0, o));
            mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, o));
            if (NCH == 2) {
                mn1 = fminf(mn1, __shfl_xor_sync(0xffffffffu, mn1, o));
                mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, o));
            }
        }
        if (lane == 0) {
            set_float(mn0, mx0, out + (size_t) i * 2);
            if (NCH == 2) set_float(mn1, mx1, out + ((size_t) nThumb + i) * 2);
        }
    }
}

// Pixel boundaries of CachedWindow::refillCache (juce_AudioThumbnail.cpp:459-518): the reference keeps a
// RUNNING sum startTime += timePerPixel in double, so the boundaries are produced by one thread in the
// same order.  bnd[i] = first index of pixel i, bnd[i+1] = nextSample.  factor = rate (zoomed-in branch)
// or rate / samplesPerThumbSample.
__global__ void thumb_bounds_kernel(double startTime, double timePerPixel, double factor, int width, int* bnd) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    int sample = juce_round_to_int(__dmul_rn(startTime, factor));
    bnd[0] = sample;
    for (int i = 0; i < width; ++i) {
        const int next = juce_round_to_int(__dmul_rn(__dadd_rn(startTime, timePerPixel), factor));
        bnd[i + 1] = next;
        startTime = __dadd_rn(startTime, timePerPixel);
        sample = next;
    }
}

// Zoomed-out branch (K7): ThumbData::getMinMax over the level-0 pairs, inclusive range
// (juce_AudioThumbnail.cpp:292-319, 493-518).  One thread per (channel, pixel).
__global__ void __launch_bounds__(256)
thumb_pixels_kernel(const int8_t* __restrict__ lvl0, int nThumb, int nch, const int* __restrict__ bnd, int width,
                    int8_t* __restrict__ out) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= nch * width) return;
    const int ch = idx / width, i = idx - ch * width;
    int s = bnd[i], e = bnd[i + 1];
    int8_t o0 = 1, o1 = 0;
    if (s >= 0) {
        e = min(e, nThumb - 1);
        int mx = -128, mn = 127;
        const char2* p = reinterpret_cast<const char2*>(lvl0) + (size_t) ch * nThumb;
        for (; s <= e; ++s) {
            const char2 v = p[s];
            mn = min(mn, (int) v.x);
            mx = max(mx, (int) v.y);
        }
        if (mn <= mx) { o0 = (int8_t) mn; o1 = (int8_t) mx; }
    }
    out[(size_t) idx * 2] = o0;
    out[(size_t) idx * 2 + 1] = o1;
}

// Zoomed-in branch: per pixel readMaxLevels(sample, max(1, next - sample)) straight from the PCM
// (juce_AudioThumbnail.cpp:459-492); frames past the clip end read as 0 (AudioFormatReader::read).
// One warp per pixel, both channels.
__global__ void __launch_bounds__(256)
thumb_pixels_direct_kernel(const float* __restrict__ data, int nch, int length, const int* __restrict__ bnd, int width,
                           int8_t* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (i >= width) return;
    const int s = bnd[i];
    if (s < 0) return;                      // reference leaves the cache entry untouched; our buffer is zero-filled
    if (s >= length) {
        if (lane == 0) for (int c = 0; c < nch; ++c) { out[((size_t) c * width + i) * 2] = 0; out[((size_t) c * width + i) * 2 + 1] = 0; }
        return;
    }
    const int num = max(1, bnd[i + 1] - s);
    float mn0 = INFINITY, mx0 = -INFINITY, mn1 = INFINITY, mx1 = -INFINITY;
    for (int k = lane; k < num; k += 32) {
        const long long f = (long long) s + k;
        float a = 0.0f, b = 0.0f;
        if (f < length) {
            if (nch > 1) { const float2 v = __ldg(reinterpret_cast<const float2*>(data) + f); a = v.x; b = v.y; }
            else a = __ldg(data + f);
        }
        mn0 = fminf(mn0, a); mx0 = fmaxf(mx0, a);
        mn1 = fminf(mn1, b); mx1 = fmaxf(mx1, b);
    }
    for (int o = 16; o > 0; o >>= 1) {
        mn0 = fminf(mn0, __shfl_xor_sync(0xffffffffu, mn0, o)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, o));
        mn1 = fminf(mn1, __shfl_xor_sync(0xffffffffu, mn1, o)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, o));
    }
    if (lane == 0) {
        set_float(mn0, mx0, out + (size_t) i * 2);
        if (nch > 1) set_float(mn1, mx1, out + ((size_t) width + i) * 2);
    }
}

// Column fill (K8): CachedWindow::drawChannel geometry (juce_AudioThumbnail.cpp:393-417) with the int
// truncation QPainter::fillRect(int,int,int,int) applies (lib/QPainterContext.cpp:190-200).
// One thread per column; mask[height][width] is pre-zeroed.
__global__ void __launch_bounds__(256)
thumb_raster_kernel(const int8_t* __restrict__ px, int width, int height, float zoom, uint8_t* __restrict__ mask) {
    const int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= width) return;
    const int lo = px[(size_t) w * 2], hi = px[(size_t) w * 2 + 1];
    if (!(hi > lo)) return;                                         // isNonZero
    const float topY = 0.0f, bottomY = (float) height;
    const float midY = __fmul_rn(__fadd_rn(topY, bottomY), 0.5f);
    const float vscale = __fdiv_rn(__fmul_rn(zoom, __fsub_rn(bottomY, topY)), 256.0f);
    const float top = fmaxf(__fsub_rn(__fsub_rn(midY, __fmul_rn((float) hi, vscale)), 0.3f), topY);
    const float bottom = fminf(__fadd_rn(__fsub_rn(midY, __fmul_rn((float) lo, vscale)), 0.3f), bottomY);
    const float x = (float) w;                                      // x += 1.0f from 0 is exact below 2^24
    const int rx = (int) x, ry = (int) top, rh = (int) __fsub_rn(bottom, top);
    if (rh > 0 && rx >= 0 && rx < width)
        for (int y = max(ry, 0); y < min(ry + rh, height); ++y) mask[(size_t) y * width + rx] = 1;
}

// Standalone JackPassthroughPrivate::process (lib/JackPassthrough.cpp:45-115) over device buffers:
// out6 = dryL,dryR,fx1L,fx1R,fx2L,fx2R, each n floats.
__global__ void __launch_bounds__(256)
passthrough_kernel(zl_passthrough_params p, const float* __restrict__ inL, const float* __restrict__ inR, long long n,
                   float* __restrict__ out6) {
    const long long stride = (long long) gridDim.x * blockDim.x;
    const float pl = fminf(__fsub_rn(1.0f, p.pan_amount), 1.0f), pr = fminf(__fadd_rn(1.0f, p.pan_amount), 1.0f);
    const float amt[3] = {p.dry_amount, p.wet_fx1_amount, p.wet_fx2_amount};
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float l = inL[i], r = inR[i];
        #pragma unroll
        for (int k = 0; k < 3; ++k) {
            float ol, orr;
            if (p.muted || (p.pan_amount == 0.0f && amt[k] == 0.0f)) { ol = 0.0f; orr = 0.0f; }
            else if (p.pan_amount == 0.0f && amt[k] == 1.0f) { ol = l; orr = r; }
            else { ol = __fmul_rn(__fmul_rn(amt[k], l), pl); orr = __fmul_rn(__fmul_rn(amt[k], r), pr); }
            out6[(2 * k) * n + i] = ol;
            out6[(2 * k + 1) * n + i] = orr;
        }
    }
}

// Standalone peak scan + sum of squares of one period (lib/AudioLevels.cpp:365-382;
// JUCE/juce_audio_basics/buffers/juce_AudioSampleBuffer.h:1087-1095).  res: peaks[2] int, sumsq[2] double
// (pre-zeroed).  Peaks are exact (max); sums are reduced per CTA then atomically (order-free in f64).
__global__ void __launch_bounds__(256)
levels_kernel(const float* __restrict__ l, const float* __restrict__ r, long long n, int* peaks, double* sumsq) {
    __shared__ double sq[8][2];
    __shared__ int pk[8][2];
    const long long stride = (long long) gridDim.x * blockDim.x;
    int pa = 0, pb = 0; double sa = 0.0, sb = 0.0;
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float a = l[i], b = r[i];
        pa = max(pa, __float2int_rz(fabsf(__fmul_rn(131072.0f, a))));
        pb = max(pb, __float2int_rz(fabsf(__fmul_rn(131072.0f, b))));
        sa += (double) __fmul_rn(a, a);
        sb += (double) __fmul_rn(b, b);
    }
    pa = __reduce_max_sync(0xffffffffu, pa); pb = __reduce_max_sync(0xffffffffu, pb);
    for (int o = 16; o > 0; o >>= 1) { sa += __shfl_down_sync(0xffffffffu, sa, o); sb += __shfl_down_sync(0xffffffffu, sb, o); }
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) { sq[w][0] = sa; sq[w][1] = sb; pk[w][0] = pa; pk[w][1] = pb; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < 8; ++k) { sa += sq[k][0]; sb += sq[k][1]; pa = max(pa, pk[k][0]); pb = max(pb, pk[k][1]); }
        atomicMax(peaks, pa); atomicMax(peaks + 1, pb);
        atomicAdd(sumsq, sa); atomicAdd(sumsq + 1, sb);
    }
}

// planar host layout -> interleaved frames (SamplerSynthSound::loadSoundData keeps min(2, nch) channels,
// lib/SamplerSynthSound.cpp:45-47), padded with zero frames.
__global__ void __launch_bounds__(256)
interleave_kernel(const float* __restrict__ ch0, const float* __restrict__ ch1, long long len, long long padded,
                  float* __restrict__ out) {
    const long long stride = (long long) gridDim.x * blockDim.x;
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < padded; i += stride) {
        if (ch1) reinterpret_cast<float2*>(out)[i] = i < len ? make_float2(ch0[i], ch1[i]) : make_float2(0.f, 0.f);
        else out[i] = i < len ? ch0[i] : 0.0f;
    }
}

// ---- launch wrappers (called from zl_api.cu) -------------------------------------------------------
static inline int grid_for(long long n, int threads, int cap) {
    long long g = (n + threads - 1) / threads;
    if (g < 1) g = 1;
    return (int) (g < cap ? g : cap);
}

cudaError_t launch_thumb_level0(const float* data, int nch, int spt, int nDone, int nThumb, int8_t* out, cudaStream_t st) {
    if (nDone <= 0) return cudaSuccess;
    const int grid = grid_for((long long) nDone * 32, 256, 148 * 64);
    if (nch > 1) thumb_level0_kernel<2><<<grid, 256, 0, st>>>(data, spt, nDone, nThumb, out);
    else thumb_level0_kernel<1><<<grid, 256, 0, st>>>(data, spt, nDone, nThumb, out);
    return cudaGetLastError();
}
cudaError_t launch_thumb_bounds(double t0, double dt, double factor, int width, int* bnd, cudaStream_t st) {
    thumb_bounds_kernel<<<1, 32, 0, st>>>(t0, dt, factor, width, bnd);
    return cudaGetLastError();
}
cudaError_t launch_thumb_pixels(const int8_t* lvl0, int nThumb, int nch, const int* bnd, int width, int8_t* out, cudaStream_t st) {
    thumb_pixels_kernel<<<grid_for((long long) nch * width, 256, 1 << 30), 256, 0, st>>>(lvl0, nThumb, nch, bnd, width, out);
    return cudaGetLastError();
}
cudaError_t launch_thumb_pixels_direct(const float* data, int nch, int length, const int* bnd, int width, int8_t* out, cudaStream_t st) {
    thumb_pixels_direct_kernel<<<grid_for((long long) width * 32, 256, 1 << 30), 256, 0, st>>>(data, nch, length, bnd, width, out);
    return cudaGetLastError();
}
cudaError_t launch_thumb_raster(const int8_t* px, int width, int height, float zoom, uint8_t* mask, cudaStream_t st) {
    thumb_raster_kernel<<<grid_for(width, 256, 1 << 30), 256, 0, st>>>(px, width, height, zoom, mask);
    return cudaGetLastError();
}
cudaError_t launch_passthrough(const zl_passthrough_params& p, const float* l, const float* r, long long n, float* out6, cudaStream_t st) {
    passthrough_kernel<<<grid_for(n, 256, 148 * 8), 256, 0, st>>>(p, l, r, n, out6);
    return cudaGetLastError();
}
cudaError_t launch_levels(const float* l, const float* r, long long n, int* peaks, double* sumsq, cudaStream_t st) {
    levels_kernel<<<grid_for(n, 256, 148 * 8), 256, 0, st>>>(l, r, n, peaks, sumsq);
    return cudaGetLastError();
}
cudaError_t launch_interleave(const float* c0, const float* c1, long long len, long long padded, float* out, cudaStream_t st) {
    interleave_kernel<<<grid_for(padded, 256, 148 * 16), 256, 0, st>>>(c0, c1, len, padded, out);
    return cudaGetLastError();
}

}  // namespace zl
