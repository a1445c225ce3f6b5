// zl_wav.cu — rows n1 and n3 of SURVEY.md §8f: the data formats either side of the render path.
//
//   n1  WAV image -> HBM clip.  The RIFF/WAVE chunk walk of juce::WavAudioFormatReader
//       (JUCE/juce_audio_formats/codecs/juce_WavAudioFormat.cpp:933-1190) runs on the host over the caller's bytes; the
//       sample conversion — ReadHelper<Int32, {UInt8,Int16,Int24,Int32}> left-justifies every integer sample into an
//       int32 (juce_AudioDataConverters.h:128-225), AudioFormatReader::read then scales it with
//       f32(i) * (1.0f / 0x7fffffff) (juce_AudioFormatReader.cpp:39-46,139-183; the factor rounds to exactly 2^-31) —
//       runs on the device, straight from the file's interleaved bytes to the interleaved f32 frames the gather reads.
//       Channel selection is SamplerSynthSound::loadSoundData's: AudioBuffer<float>(min(2, nch), length), read(...,
//       true, true) (lib/SamplerSynthSound.cpp:40-46) = the file's first two channels.
//   n3  device plane -> 16/24-bit WAV.  AudioFormatWriter::writeFromFloatArrays turns floats into int32 with
//       convertFloatsToInts (juce_AudioFormatWriter.cpp:64-79: <= -1 -> INT_MIN, >= 1 -> INT_MAX, else
//       roundToInt(INT_MAX * (double) x)), WavAudioFormatWriter::write keeps the top 16 / 24 bits (juce_WavAudioFormat.cpp
//       :1336-1369, juce_AudioDataConverters.h:159,184); the header is the writer's RIFF | JUNK(52) | fmt (16) | data
//       (:1400-1500) so that the bytes equal what the reference's DiskWriter (lib/AudioLevels.cpp:35-119) leaves on disk.
//
// Both are HBM-bound byte kernels: every input byte is read once (16-byte vector loads staged through shared memory,
// because 24-bit frames are not naturally aligned) and every output element written once, fully coalesced.
#include <cstdint>
#include <cstring>

#include <cuda_runtime.h>

#include "../../include/zl_b200.h"

namespace zl {

// ---- host: chunk walk ---------------------------------------------------------------------------------------------
namespace {
struct Cursor {
    const unsigned char* p; uint64_t n, pos;
    bool exhausted() const { return pos >= n; }
    // juce::InputStream::readInt / readShort: short reads return zero-filled values
    uint32_t u32() { unsigned char b[4] = {0, 0, 0, 0}; for (int i = 0; i < 4; ++i) if (pos + i < n) b[i] = p[pos + i]; pos += 4; return b[0] | (b[1] << 8) | (b[2] << 16) | ((uint32_t) b[3] << 24); }
    uint16_t u16() { unsigned char b[2] = {0, 0}; for (int i = 0; i < 2; ++i) if (pos + i < n) b[i] = p[pos + i]; pos += 2; return (uint16_t) (b[0] | (b[1] << 8)); }
    uint64_t u64() { const uint64_t lo = u32(); const uint64_t hi = u32(); return lo | (hi << 32); }
};
constexpr uint32_t tag(const char (&s)[5]) { return (uint32_t) (unsigned char) s[0] | ((uint32_t) (unsigned char) s[1] << 8) | ((uint32_t) (unsigned char) s[2] << 16) | ((uint32_t) (unsigned char) s[3] << 24); }
const unsigned char kPcmGuidTail[14] = {0x00, 0x00, 0x10, 0x00, 0x80, 0x00, 0x00, 0xaa, 0x00, 0x38, 0x9b, 0x71, 0, 0};
}  // namespace

// Fills `f`; returns ZL_OK when WavAudioFormat::createReaderFor would hand out a reader (sampleRate > 0, numChannels > 0,
// bytesPerFrame > 0, bitsPerSample <= 32; juce_WavAudioFormat.cpp:1718-1731) whose samples copySampleData can decode.
int wav_parse(const void* bytes, uint64_t n, zl_wav_format* f) {
    if (!bytes || !f) return ZL_ERR_INVALID;
    memset(f, 0, sizeof *f);
    Cursor in{static_cast<const unsigned char*>(bytes), n, 0};
    uint64_t len = 0, end = 0, dataLength = 0;
    bool rf64 = false;
    const uint32_t first = in.u32();
    if (first == tag("RF64")) { in.pos += 4; rf64 = true; }                     // :947-951
    else if (first == tag("RIFF")) { len = in.u32(); end = len + in.pos; }      // :952-956
    else return ZL_ERR_INVALID;
    const uint64_t startOfRiff = in.pos;
    if (in.u32() != tag("WAVE")) return ZL_ERR_INVALID;
    if (rf64 && in.u32() == tag("ds64")) {                                       // :965-977
        const uint32_t length = in.u32();
        if (length < 28) return ZL_ERR_INVALID;
        const uint64_t chunkEnd = in.pos + length + (length & 1);
        len = in.u64(); end = len + startOfRiff; dataLength = in.u64();
        in.pos = chunkEnd;
    }
    int bytesPerFrame = 0;
    while (in.pos < end && !in.exhausted()) {                                    // :979-1183
        const uint64_t chunkHeader = in.pos;
        const uint32_t type = in.u32();
        const uint32_t length = in.u32();
        uint64_t chunkEnd = in.pos + length + (length & 1);
        if (type == tag("fmt ")) {
            const uint16_t format = in.u16();
            f->channels = (int32_t) (int16_t) in.u16();
            f->sample_rate = (double) (int32_t) in.u32();
            const int32_t bytesPerSec = (int32_t) in.u32();
            in.pos += 2;
            uint32_t bits = (uint32_t) (int32_t) (int16_t) in.u16();
            if (bits > 64 && (int) f->sample_rate != 0) {
                bytesPerFrame = bytesPerSec / (int) f->sample_rate;
                if (f->channels != 0) bits = 8 * (uint32_t) bytesPerFrame / (uint32_t) f->channels;
            } else bytesPerFrame = (int) ((uint32_t) f->channels * bits / 8);
            f->is_float = 0;
            if (format == 3) f->is_float = 1;
            else if (format == 0xfffe) {                                         // WAVE_FORMAT_EXTENSIBLE :1012-1037
                if (length < 40) bytesPerFrame = 0;
                else {
                    in.pos += 4 + 4;                                             // cbSize + valid bits, channel mask
                    const uint32_t d1 = in.u32();
                    unsigned char tail[12];
                    for (int i = 0; i < 12; ++i) { tail[i] = in.pos < in.n ? in.p[in.pos] : 0; in.pos++; }
                    const bool std_tail = memcmp(tail, kPcmGuidTail, 12) == 0;   // data2..data4 of the KSDATAFORMAT GUIDs
                    const unsigned char ambi[12] = {0x21, 0x07, 0xd3, 0x11, 0x86, 0x44, 0xc8, 0xc1, 0xca, 0x00, 0x00, 0x00};
                    if (d1 == 3 && std_tail) f->is_float = 1;
                    else if (!((d1 == 1 && std_tail) || (d1 == 1 && memcmp(tail, ambi, 12) == 0))) bytesPerFrame = 0;
                }
            } else if (format >= 0x674f && format <= 0x6771 && (format <= 0x6751 || format >= 0x676f)) return ZL_ERR_INVALID;   // Ogg in WAV
            else if (format != 1) bytesPerFrame = 0;
            f->bits_per_sample = (int32_t) bits;
        } else if (type == tag("data")) {
            if (rf64) { if (dataLength > 0) chunkEnd = in.pos + dataLength + (dataLength & 1); }
            else dataLength = length;
            f->data_offset = in.pos;
            f->data_bytes = dataLength;
            f->frames = bytesPerFrame > 0 ? (int64_t) (dataLength / (uint64_t) bytesPerFrame) : 0;
        } else if (type != tag("bext") && type != tag("smpl") && type != tag("inst") && type != tag("INST") && type != tag("cue ")
                   && type != tag("axml") && type != tag("LIST") && type != tag("acid") && type != tag("Trkn") && chunkEnd <= in.pos) break;   // :1176-1179
        // a ds64 data length near 2^64 wraps chunkEnd back into the image (the reference's int64 sum does the same and its walk
        // never ends on such a file): the walk stops unless the chunk ends after its own header
        if (chunkEnd <= chunkHeader) break;
        in.pos = chunkEnd;
    }
    f->bytes_per_frame = bytesPerFrame;
    if (!(f->sample_rate > 0) || f->channels <= 0 || bytesPerFrame <= 0 || f->bits_per_sample > 32) return ZL_ERR_INVALID;
    const int b = f->bits_per_sample;
    if (!(b == 8 || b == 16 || b == 24 || b == 32) || (f->is_float && b != 32)) return ZL_ERR_INVALID;   // copySampleData's cases :1232-1245
    if (f->channels * (b / 8) > bytesPerFrame) return ZL_ERR_INVALID;
    return ZL_OK;
}

// ---- device: decode -----------------------------------------------------------------------------------------------
enum { kU8 = 0, kS16 = 1, kS24 = 2, kS32 = 3, kF32 = 4 };

template <int FMT>
__device__ __forceinline__ float wav_sample(const unsigned char* q) {
    int32_t i;
    if (FMT == kU8) i = (int32_t) (((uint32_t) q[0] - 128u) << 24);                            // juce_AudioDataConverters.h:134
    else if (FMT == kS16) i = (int32_t) (((uint32_t) q[0] | ((uint32_t) q[1] << 8)) << 16);    // :159
    else if (FMT == kS24) i = (int32_t) (((uint32_t) q[0] | ((uint32_t) q[1] << 8) | ((uint32_t) q[2] << 16)) << 8);   // :184
    else i = (int32_t) ((uint32_t) q[0] | ((uint32_t) q[1] << 8) | ((uint32_t) q[2] << 16) | ((uint32_t) q[3] << 24));
    if (FMT == kF32) return __int_as_float(i);                                                  // bits pass through
    return __fmul_rn(__int2float_rn(i), 4.656612873077393e-10f);                                // f32(i) * 2^-31
}

// CTA = framesPerCta consecutive frames.  raw: the data chunk, 16-byte aligned, readable (zero-filled) up to the next
// multiple of 16 past rawBytes.  out: [padded][nchOut] interleaved; frames past `frames` are the zero padding.
template <int FMT>
__global__ void __launch_bounds__(256) wav_decode_kernel(const unsigned char* __restrict__ raw, int bytesPerFrame, int nchOut,
                                                         long long frames, long long padded, int framesPerCta, float* __restrict__ out) {
    constexpr int BPS = FMT == kU8 ? 1 : (FMT == kS16 ? 2 : (FMT == kS24 ? 3 : 4));
    extern __shared__ uint4 stage[];
    const long long f0 = (long long) blockIdx.x * framesPerCta;
    const int nf = (int) min((long long) framesPerCta, padded - f0);
    const int nfSrc = (int) max(0LL, min((long long) nf, frames - f0));
    const unsigned long long b0 = (unsigned long long) f0 * bytesPerFrame, b1 = b0 + (unsigned long long) nfSrc * bytesPerFrame;
    const unsigned long long a0 = b0 & ~15ull;
    const int nvec = nfSrc ? (int) ((b1 - a0 + 15) >> 4) : 0;
    const uint4* src = reinterpret_cast<const uint4*>(raw + a0);
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) stage[i] = __ldcs(src + i);          // read once: streaming
    __syncthreads();
    const unsigned char* s = reinterpret_cast<const unsigned char*>(stage) + (b0 - a0);
    float* o = out + f0 * nchOut;
    for (int k = threadIdx.x; k < nf * nchOut; k += blockDim.x) {
        const int f = nchOut == 2 ? k >> 1 : k, c = nchOut == 2 ? k & 1 : 0;
        o[k] = f < nfSrc ? wav_sample<FMT>(s + (size_t) f * bytesPerFrame + c * BPS) : 0.0f;
    }
}

// The layouts nearly every file has - 16-bit or 32-bit samples, all of the file's channels kept - are one flat array of
// samples: thread = 16 input bytes (8 x s16 or 4 x 32-bit) -> 8 or 4 floats, vector loads and stores, no staging.
// total = samples to decode, outTotal = floats to write (the rest is the zero padding).
template <int FMT>
__global__ void __launch_bounds__(256) wav_decode_flat_kernel(const uint4* __restrict__ raw, long long total, long long outTotal, float* __restrict__ out) {
    constexpr int PER = FMT == kS16 ? 8 : 4;
    const long long v = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    const long long k0 = v * PER;
    if (k0 >= outTotal) return;
    float f[PER];
    if (k0 < total) {                                      // raw is readable (zero-filled) to the end of this vector
        const uint4 w = __ldcs(raw + v);
        const uint32_t u[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (FMT == kS16) {
                f[2 * i] = __fmul_rn(__int2float_rn((int32_t) (u[i] << 16)), 4.656612873077393e-10f);
                f[2 * i + 1] = __fmul_rn(__int2float_rn((int32_t) (u[i] & 0xffff0000u)), 4.656612873077393e-10f);
            } else if (FMT == kS32) f[i] = __fmul_rn(__int2float_rn((int32_t) u[i]), 4.656612873077393e-10f);
            else f[i] = __uint_as_float(u[i]);
        }
#pragma unroll
        for (int i = 0; i < PER; ++i) if (k0 + i >= total) f[i] = 0.0f;
    } else {
#pragma unroll
        for (int i = 0; i < PER; ++i) f[i] = 0.0f;
    }
    if (k0 + PER <= outTotal) {
#pragma unroll
        for (int i = 0; i < PER; i += 4) __stcs(reinterpret_cast<float4*>(out + k0 + i), make_float4(f[i], f[i + 1], f[i + 2], f[i + 3]));
    } else {
        for (int i = 0; i < PER && k0 + i < outTotal; ++i) out[k0 + i] = f[i];
    }
}

cudaError_t launch_wav_decode(const unsigned char* raw, const zl_wav_format& f, int nchOut, long long padded, float* out, cudaStream_t st) {
    const int fmt = f.is_float ? kF32 : (f.bits_per_sample == 8 ? kU8 : (f.bits_per_sample == 16 ? kS16 : (f.bits_per_sample == 24 ? kS24 : kS32)));
    const int bps = f.bits_per_sample / 8;
    if ((fmt == kS16 || fmt == kS32 || fmt == kF32) && f.bytes_per_frame == nchOut * bps) {
        const long long total = (long long) f.frames * nchOut, outTotal = padded * nchOut;
        const int per = fmt == kS16 ? 8 : 4;
        const unsigned g = (unsigned) (((outTotal + per - 1) / per + 255) / 256);
        const uint4* r4 = reinterpret_cast<const uint4*>(raw);
        if (fmt == kS16) wav_decode_flat_kernel<kS16><<<g, 256, 0, st>>>(r4, total, outTotal, out);
        else if (fmt == kS32) wav_decode_flat_kernel<kS32><<<g, 256, 0, st>>>(r4, total, outTotal, out);
        else wav_decode_flat_kernel<kF32><<<g, 256, 0, st>>>(r4, total, outTotal, out);
        return cudaGetLastError();
    }
    int framesPerCta = 40960 / f.bytes_per_frame;
    if (framesPerCta > 2048) framesPerCta = 2048;
    if (framesPerCta < 1) return cudaErrorInvalidValue;                          // > 40 KB per frame: not a sampler clip
    const size_t smem = (size_t) framesPerCta * f.bytes_per_frame + 32;
    const unsigned grid = (unsigned) ((padded + framesPerCta - 1) / framesPerCta);
#define ZL_WAV_DECODE(F) wav_decode_kernel<F><<<grid, 256, smem, st>>>(raw, f.bytes_per_frame, nchOut, (long long) f.frames, padded, framesPerCta, out)
    switch (fmt) {
        case kU8: ZL_WAV_DECODE(kU8); break;
        case kS16: ZL_WAV_DECODE(kS16); break;
        case kS24: ZL_WAV_DECODE(kS24); break;
        case kS32: ZL_WAV_DECODE(kS32); break;
        default: ZL_WAV_DECODE(kF32); break;
    }
#undef ZL_WAV_DECODE
    return cudaGetLastError();
}

// clip frames [len][nch] -> planar [nch][len] (zl_clip_read_pcm)
__global__ void __launch_bounds__(256) deinterleave_kernel(const float* __restrict__ in, int nch, long long len, float* __restrict__ out) {
    const long long j = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= len) return;
    if (nch == 2) { const float2 v = __ldcs(reinterpret_cast<const float2*>(in) + j); out[j] = v.x; out[len + j] = v.y; }
    else out[j] = __ldcs(in + j);
}
cudaError_t launch_deinterleave(const float* in, int nch, long long len, float* out, cudaStream_t st) {
    deinterleave_kernel<<<(unsigned) ((len + 255) / 256), 256, 0, st>>>(in, nch, len, out);
    return cudaGetLastError();
}

// n4: level-0 table [nch][n](min,max) <-> jatm order [n][nch](min,max); one 2-byte pair per thread element, nch <= 2
__global__ void __launch_bounds__(256) jatm_transpose_kernel(const uint16_t* __restrict__ in, int nch, int n, int toJatm, uint16_t* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    for (int ch = 0; ch < nch; ++ch) {
        if (toJatm) out[(size_t) i * nch + ch] = in[(size_t) ch * n + i];
        else out[(size_t) ch * n + i] = in[(size_t) i * nch + ch];
    }
}
cudaError_t launch_jatm_transpose(const void* in, int nch, int n, int toJatm, void* out, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    jatm_transpose_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, st>>>(static_cast<const uint16_t*>(in), nch, n, toJatm, static_cast<uint16_t*>(out));
    return cudaGetLastError();
}

// ---- device: encode -----------------------------------------------------------------------------------------------
// convertFloatsToInts (juce_AudioFormatWriter.cpp:64-79); roundToInt of a double = round half to even (juce_MathsFunctions.h:471-485)
__device__ __forceinline__ int32_t float_to_fixed(float x) {
    const double s = (double) x;
    if (s <= -1.0) return INT32_MIN;
    if (s >= 1.0) return INT32_MAX;
    return __double2int_rn(__dmul_rn(2147483647.0, s));
}

// l, r: planar rows of n frames (r == nullptr: mono file).  out: interleaved little-endian samples of BYTES bytes each.
template <int BYTES>
__global__ void __launch_bounds__(256) wav_encode_kernel(const float* __restrict__ l, const float* __restrict__ r, long long n, unsigned char* __restrict__ out) {
    const long long j = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const int nch = r ? 2 : 1;
    unsigned char* o = out + (size_t) j * nch * BYTES;
    for (int c = 0; c < nch; ++c) {
        const int32_t v = float_to_fixed(__ldcs((c ? r : l) + j));
        if (BYTES == 2) {                                                       // Int16::setAsInt32LE: v >> 16   :159
            const uint16_t h = (uint16_t) (v >> 16);
            if (nch == 2) { reinterpret_cast<uint16_t*>(o)[c] = h; } else *reinterpret_cast<uint16_t*>(o) = h;
        } else {                                                                // Int24::setAsInt32LE: v >> 8    :184
            const int32_t t = v >> 8;
            o[c * 3] = (unsigned char) t; o[c * 3 + 1] = (unsigned char) (t >> 8); o[c * 3 + 2] = (unsigned char) (t >> 16);
        }
    }
}

// Stereo, rows 16-byte aligned: thread = 4 frames -> 16 (16-bit) or 24 (24-bit) packed bytes, vector loads and stores.
template <int BYTES>
__device__ __forceinline__ void encode4(const float4 a, const float4 b, unsigned char* __restrict__ out, long long j) {
    const int32_t v[8] = {float_to_fixed(a.x), float_to_fixed(b.x), float_to_fixed(a.y), float_to_fixed(b.y),
                          float_to_fixed(a.z), float_to_fixed(b.z), float_to_fixed(a.w), float_to_fixed(b.w)};
    if (BYTES == 2) {
        uint4 w;
        w.x = ((uint32_t) v[0] >> 16) | ((uint32_t) v[1] & 0xffff0000u); w.y = ((uint32_t) v[2] >> 16) | ((uint32_t) v[3] & 0xffff0000u);
        w.z = ((uint32_t) v[4] >> 16) | ((uint32_t) v[5] & 0xffff0000u); w.w = ((uint32_t) v[6] >> 16) | ((uint32_t) v[7] & 0xffff0000u);
        __stcs(reinterpret_cast<uint4*>(out) + j, w);
    } else {
        uint32_t t[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) t[i] = (uint32_t) v[i] >> 8;                        // 24 significant bits, little endian
        uint2 w0, w1, w2;
        w0.x = t[0] | (t[1] << 24); w0.y = (t[1] >> 8) | (t[2] << 16);
        w1.x = (t[2] >> 16) | (t[3] << 8); w1.y = t[4] | (t[5] << 24);
        w2.x = (t[5] >> 8) | (t[6] << 16); w2.y = (t[6] >> 16) | (t[7] << 8);
        uint2* o = reinterpret_cast<uint2*>(out) + j * 3;
        __stcs(o, w0); __stcs(o + 1, w1); __stcs(o + 2, w2);
    }
}
template <int BYTES>
__global__ void __launch_bounds__(256) wav_encode4_kernel(const float4* __restrict__ l, const float4* __restrict__ r, long long n4, unsigned char* __restrict__ out) {
    const long long j = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n4) return;
    encode4<BYTES>(__ldcs(l + j), __ldcs(r + j), out, j);
}

// Every session's master pair at once (the batched bounce): rows [session][L|R][stride] -> [session][n frames][L,R] samples of
// BYTES bytes, i.e. the data chunk of each session's file back to back.  blockIdx.y = session.  VEC: n, stride multiples of 4.
template <int BYTES, bool VEC>
__global__ void __launch_bounds__(256) master_pcm_kernel(const float* __restrict__ master, long long stride, long long n, unsigned char* __restrict__ out) {
    const long long j = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    const float* l = master + (size_t) blockIdx.y * 2 * stride;
    unsigned char* o = out + (size_t) blockIdx.y * n * 2 * BYTES;
    if (VEC) {
        if (j >= n / 4) return;
        encode4<BYTES>(__ldcs(reinterpret_cast<const float4*>(l) + j), __ldcs(reinterpret_cast<const float4*>(l + stride) + j), o, j);
    } else {
        if (j >= n) return;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const int32_t v = float_to_fixed(__ldcs(l + c * stride + j));
            unsigned char* q = o + ((size_t) j * 2 + c) * BYTES;
            if (BYTES == 2) *reinterpret_cast<uint16_t*>(q) = (uint16_t) (v >> 16);
            else { const int32_t t = v >> 8; q[0] = (unsigned char) t; q[1] = (unsigned char) (t >> 8); q[2] = (unsigned char) (t >> 16); }
        }
    }
}
cudaError_t launch_master_pcm(const float* master, long long stride, int nSessions, long long n, int bits, unsigned char* out, cudaStream_t st) {
    if (n <= 0 || nSessions <= 0) return cudaSuccess;
    if (bits != 16 && bits != 24) return cudaErrorInvalidValue;
    const bool vec = (n % 4) == 0 && (stride % 4) == 0 && (reinterpret_cast<uintptr_t>(master) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0;
    const dim3 grid((unsigned) (((vec ? n / 4 : n) + 255) / 256), (unsigned) nSessions);
    if (vec) { if (bits == 16) master_pcm_kernel<2, true><<<grid, 256, 0, st>>>(master, stride, n, out); else master_pcm_kernel<3, true><<<grid, 256, 0, st>>>(master, stride, n, out); }
    else { if (bits == 16) master_pcm_kernel<2, false><<<grid, 256, 0, st>>>(master, stride, n, out); else master_pcm_kernel<3, false><<<grid, 256, 0, st>>>(master, stride, n, out); }
    return cudaGetLastError();
}

cudaError_t launch_wav_encode(const float* l, const float* r, long long n, int bits, unsigned char* out, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    if (bits != 16 && bits != 24) return cudaErrorInvalidValue;
    long long done = 0;
    if (r && ((reinterpret_cast<uintptr_t>(l) | reinterpret_cast<uintptr_t>(r)) & 15) == 0 && n >= 4) {
        const long long n4 = n / 4;
        const unsigned g4 = (unsigned) ((n4 + 255) / 256);
        if (bits == 16) wav_encode4_kernel<2><<<g4, 256, 0, st>>>(reinterpret_cast<const float4*>(l), reinterpret_cast<const float4*>(r), n4, out);
        else wav_encode4_kernel<3><<<g4, 256, 0, st>>>(reinterpret_cast<const float4*>(l), reinterpret_cast<const float4*>(r), n4, out);
        done = n4 * 4;
        if (done == n) return cudaGetLastError();
        l += done; r += done; out += (size_t) done * 2 * (bits / 8); n -= done;      // the last 1-3 frames: scalar kernel
    }
    const unsigned grid = (unsigned) ((n + 255) / 256);
    if (bits == 16) wav_encode_kernel<2><<<grid, 256, 0, st>>>(l, r, n, out);
    else if (bits == 24) wav_encode_kernel<3><<<grid, 256, 0, st>>>(l, r, n, out);
    else return cudaErrorInvalidValue;
    return cudaGetLastError();
}

// WavAudioFormatWriter::writeHeader for a file without metadata, below 4 GB: RIFF | WAVE | JUNK (52 zero bytes, room for a
// later ds64) | fmt (16) | data   (juce_WavAudioFormat.cpp:1400-1500).  Returns the header size (104).
int wav_header(unsigned char* h, int nch, double rate, int bits, uint64_t dataBytes) {
    auto w32 = [&](int at, uint32_t v) { h[at] = (unsigned char) v; h[at + 1] = (unsigned char) (v >> 8); h[at + 2] = (unsigned char) (v >> 16); h[at + 3] = (unsigned char) (v >> 24); };
    auto w16 = [&](int at, uint16_t v) { h[at] = (unsigned char) v; h[at + 1] = (unsigned char) (v >> 8); };
    memset(h, 0, 104);
    const int bytesPerFrame = nch * bits / 8;
    memcpy(h, "RIFF", 4); w32(4, (uint32_t) (dataBytes + (dataBytes & 1) + 104 - 8));
    memcpy(h + 8, "WAVE", 4);
    memcpy(h + 12, "JUNK", 4); w32(16, 52);
    memcpy(h + 72, "fmt ", 4); w32(76, 16);
    w16(80, 1); w16(82, (uint16_t) nch); w32(84, (uint32_t) (int) rate); w32(88, (uint32_t) (int) ((double) bytesPerFrame * rate));
    w16(92, (uint16_t) bytesPerFrame); w16(94, (uint16_t) bits);
    memcpy(h + 96, "data", 4); w32(100, (uint32_t) dataBytes);
    return 104;
}

}  // namespace zl
