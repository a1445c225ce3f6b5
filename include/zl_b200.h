/*
 * zl_b200.h — C ABI of the B200-native sampler voice-render path of zynthbox/libzl.
 *
 * This is the drop-in boundary.  Every entry point is `extern "C"`,
 * takes plain pointers and sizes, returns an int status (ZL_OK == 0) and never throws.
 * Each group cites the reference interface it replaces (paths relative to
 * /root/reference; JUCE/ = tracktion_engine/modules/juce/modules/).
 *
 * The same header is compiled twice:
 *   - the product   (libzl_b200/csrc, CUDA sm_100a)  exports  zl_*   symbols;
 *   - the CPU oracle (oracle/, test infrastructure)    exports  zlo_*  symbols
 *     (ZL_ORACLE_BUILD).  The product never links, loads or calls the oracle.
 */
#ifndef ZL_B200_H
#define ZL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifdef ZL_ORACLE_BUILD
#define ZL_FN(name) zlo_##name
#else
#define ZL_FN(name) zl_##name
#endif
#define ZL_EXPORT __attribute__((visibility("default")))

/* ---- status codes ------------------------------------------------------------------ */
enum {
    ZL_OK = 0,
    ZL_ERR_INVALID = 1,   /* bad argument */
    ZL_ERR_CUDA = 2,      /* CUDA runtime error (latched; see zl_last_error) */
    ZL_ERR_NOMEM = 3,
    ZL_ERR_STATE = 4,     /* call order / capacity problem */
    ZL_ERR_NODEVICE = 5   /* no CUDA device: there is NO CPU fallback in the product */
};

/* ---- modes ------------------------------------------------------------------------- */
/* Z: libzl SamplerSynthVoice::process  lib/SamplerSynthVoice.cpp:174-270
 * J: juce::SamplerVoice::renderNextBlock JUCE/juce_audio_formats/sampler/juce_Sampler.cpp:117-163 */
enum { ZL_MODE_Z = 0, ZL_MODE_J = 1 };

/* Reference quirks of Z-mode.  Faithful == bit set. */
enum {
    ZL_Q1_TAP_GAIN   = 1, /* gain/env/volume multiply only the 2nd interpolation tap  (:204-205) */
    ZL_Q2_PREINC     = 2, /* frame f is accumulated at index f+1, index 0 never written (:218-221) */
    ZL_Q3_RENOTEOFF  = 4, /* noteOff() re-armed on every frame of the release window   (:253-256) */
    ZL_QUIRKS_FAITHFUL = 7
};

/* Which device buffers a render launch materialises in HBM. */
enum {
    ZL_OUT_RAW    = 1,   /* SamplerChannel left_out/right_out           lib/SamplerSynth.cpp:132-140 */
    ZL_OUT_DRY    = 2,   /* JackPassthrough dryOutLeft/Right            lib/JackPassthrough.cpp:100-101 */
    ZL_OUT_FX1    = 4,   /* JackPassthrough wetOutFx1Left/Right         :104-105 */
    ZL_OUT_FX2    = 8,   /* JackPassthrough wetOutFx2Left/Right         :108-109 */
    ZL_OUT_MASTER = 16,  /* per-session sum over buses of the dry outs (the JACK port mix) */
    ZL_OUT_LEVELS = 32,  /* per-block per-bus integer peaks              lib/AudioLevels.cpp:357-385 */
    ZL_OUT_RMS    = 64,  /* per-block per-bus sum of squares (f64)       JUCE/juce_audio_basics/buffers/juce_AudioSampleBuffer.h:1078-1096 */
    ZL_OUT_ALL    = 127
};

/* Selector for zl_read_bus. */
enum { ZL_BUS_RAW = 0, ZL_BUS_DRY = 1, ZL_BUS_FX1 = 2, ZL_BUS_FX2 = 3 };

/* ---- PODs -------------------------------------------------------------------------- */
typedef struct zl_ctx zl_ctx;

typedef struct {
    double  sample_rate;        /* output (JACK) rate: juce::Synthesiser::setCurrentPlaybackSampleRate, lib/SamplerSynth.cpp:271-272 */
    int32_t n_sessions;         /* independent mixes (one JACK graph each); 1 for the realtime layout */
    int32_t buses_per_session;  /* SamplerChannels per session; reference 12    lib/SamplerSynth.cpp:258 */
    int32_t voices_per_bus;     /* SAMPLER_CHANNEL_VOICE_COUNT; reference 8     lib/SamplerSynth.cpp:23; 1 .. 1024 */
    int32_t max_clips;
    int32_t max_slices;         /* slice table capacity per clip; reference default 16  lib/ClipAudioSource.cpp:204 */
    int32_t max_frames;         /* frames per block upper bound (JACK period) */
    int32_t max_blocks;         /* blocks per launch upper bound */
    int32_t mode;               /* ZL_MODE_Z | ZL_MODE_J */
    int32_t quirks;             /* ZL_Q* bits (Z-mode) */
    int32_t outputs;            /* ZL_OUT_* bits */
    int32_t device;             /* CUDA device ordinal */
    int32_t reserved;
} zl_config;

/* ClipAudioSource parameter block  lib/ClipAudioSource.cpp:63-82,164-168 (defaults),
 * :255-277 (start/stop position), :338-346 (volumeAbsolute), :352-365 (length), :495-528 (slices),
 * :606-700 (root note, pan, ADSR). */
typedef struct {
    float   start_s;            /* startPositionInSeconds (default 0) */
    float   length_s;           /* lengthInSeconds (default: duration) ; <0 => use duration */
    float   length_beats;       /* lengthInBeats (default -1) */
    float   pan;                /* -1..1 (default 0) */
    float   volume_abs;         /* fader position 0..1 (volumeAbsolute) */
    float   attack, decay, sustain, release;  /* juce::ADSR::Parameters; libzl ctor: 0, 0.1, 1.0, 0.05 */
    int32_t root_note;          /* default 60 */
    int32_t n_slices;           /* number of entries in slice_pos (default 16 even slices) */
    int32_t reserved;
    const double* slice_pos;    /* n_slices doubles in 0..1, or NULL for even slices */
} zl_clip_params;

/* ClipCommand  lib/ClipCommand.h:11-31 (clip pointer -> clip id). */
typedef struct {
    int32_t clip;
    int32_t midi_note;          /* default -1 */
    int32_t midi_channel;       /* -2,-1 global; 0.. tracks.  bus index = midi_channel + 2  lib/SamplerSynth.cpp:330-331 */
    int32_t start_playback;
    int32_t stop_playback;
    int32_t change_slice;
    int32_t slice;              /* default -1 */
    int32_t change_looping;
    int32_t looping;
    int32_t change_pitch;
    float   pitch_change;
    int32_t change_speed;
    float   speed_ratio;
    int32_t change_gain_db;
    float   gain_db;
    int32_t change_volume;
    float   volume;
} zl_clip_command;

/* Per-block clock: jack_get_cycle_times (lib/SamplerSynth.cpp:128) and the three SyncTimer
 * getters the voice loop reads (lib/SyncTimer.cpp:990-1009). */
typedef struct {
    uint64_t current_usecs;
    uint64_t next_usecs;
    uint64_t playhead;          /* SyncTimer::jackPlayhead()  (ticks) */
    uint64_t playhead_usecs;    /* SyncTimer::jackPlayheadUsecs() */
    uint64_t subbeat_usecs;     /* SyncTimer::jackSubbeatLengthInMicroseconds() */
} zl_block_clock;

/* JackPassthroughPrivate parameters  lib/JackPassthrough.cpp:27-31. */
typedef struct {
    float   dry_amount;         /* default 1 */
    float   wet_fx1_amount;     /* default 1 */
    float   wet_fx2_amount;     /* default 1 */
    float   pan_amount;         /* default 0 */
    int32_t muted;              /* default 0 */
} zl_passthrough_params;

/* Voice table read-back (SamplerSynthVoice::isPlaying etc., lib/SamplerSynthVoice.h:32). */
typedef struct {
    double  position;           /* sourceSamplePosition */
    double  pitch_ratio;
    float   envelope;           /* ADSR envelopeVal */
    float   gain;               /* lgain == rgain */
    float   pos_gain;           /* positions model gain   = peakGain*0.5   lib/SamplerSynthVoice.cpp:266 */
    float   pos_progress;       /* positions model progress = position/sourceSampleLength */
    int32_t playing;            /* isPlaying */
    int32_t clip;               /* clip id or -1 */
    int32_t adsr_state;         /* 0 idle,1 attack,2 decay,3 sustain,4 release */
    int32_t midi_note;
    int32_t slice;
    int32_t looping;
} zl_voice_info;

/* AudioLevels meter state for one channel  lib/AudioLevels.cpp:347-412. */
typedef struct {
    int32_t peak_a, peak_b;     /* integer peaks (bit-exact quantity) */
    float   db_a, db_b;         /* convertTodbFS */
    float   db_sum;             /* addFloat(db_a, db_b) */
    float   hold_a, hold_b;     /* playback hold signals (linear) */
    float   rms_a, rms_b;       /* extension: getRMSLevel of the scanned period */
    int32_t reserved;
} zl_levels;

/* ---- context ----------------------------------------------------------------------- */
/* Replaces initJuce()/SamplerSynth::initialize (lib/libzl.cpp:358-410, lib/SamplerSynth.cpp:254-278):
 * allocates the device-resident voice table, clip table, bus buffers. */
ZL_EXPORT int ZL_FN(ctx_create)(const zl_config* cfg, zl_ctx** out);
ZL_EXPORT int ZL_FN(ctx_destroy)(zl_ctx* ctx);
ZL_EXPORT const char* ZL_FN(last_error)(zl_ctx* ctx);
ZL_EXPORT int ZL_FN(config_default)(zl_config* cfg);
/* Use an externally created cudaStream_t (e.g. torch's current stream). NULL = context's own. */
ZL_EXPORT int ZL_FN(ctx_set_stream)(zl_ctx* ctx, void* cuda_stream);
/* Two-stream render (default on): the planner of period b + 1 runs beside the gather of period b.  0 = everything on
 * the context stream in order (used by bench.py to time gather_kernel alone). */
ZL_EXPORT int ZL_FN(ctx_set_overlap)(zl_ctx* ctx, int32_t enable);
/* A render of ONE period (the live JACK cycle) is submitted as one CUDA-graph launch - the two
 * counter resets, plan_kernel, walk_kernel, gather_kernel and the port mix captured once per command slot and replayed
 * while the kernel arguments stay the same.  On by default (environment ZL_B200_GRAPH=0 turns it off); this switch is
 * for measuring the difference. */
ZL_EXPORT int ZL_FN(ctx_set_graph)(zl_ctx* ctx, int32_t enable);
ZL_EXPORT int ZL_FN(sync)(zl_ctx* ctx);

/* ---- clips: ClipAudioSource + SamplerSynthSound ------------------------------------- */
/* Replaces ClipAudioSource_new + SamplerSynth::registerClip + SamplerSynthSound::loadSoundData
 * (lib/libzl.cpp:107-128, lib/SamplerSynth.cpp:285-295, lib/SamplerSynthSound.cpp:28-59):
 * planar host float PCM (min(2,nch) channels are kept) becomes an HBM-resident interleaved buffer. */
ZL_EXPORT int ZL_FN(clip_create_from_pcm)(zl_ctx* ctx, const float* const* planar, int32_t nch,
                                          int64_t length, double sample_rate, int32_t* clip_id);
/* Same, but the PCM is already on the device as interleaved frames (float for mono, float2 for
 * stereo); the buffer is adopted, not copied, and must outlive the clip.  It must hold length + 8 frames
 * with frames length .. length + 7 zero (the gather reads its interpolation halo and the frames of
 * lanes past a voice's end from that padding instead of predicating the loads). */
ZL_EXPORT int ZL_FN(clip_create_from_device)(zl_ctx* ctx, const void* d_interleaved, int32_t nch,
                                             int64_t length, double sample_rate, int32_t* clip_id);
/* n1: WAV ingest.  What juce::WavAudioFormatReader finds in a RIFF/WAVE (or RF64) image
 * (JUCE/juce_audio_formats/codecs/juce_WavAudioFormat.cpp:933-1190): PCM 8/16/24/32 bit or IEEE float 32. */
typedef struct zl_wav_format {
    int32_t channels;
    int32_t bits_per_sample;
    int32_t is_float;           /* usesFloatingPointData  :1008-1033 */
    int32_t bytes_per_frame;
    double sample_rate;
    int64_t frames;             /* lengthInSamples = data bytes / bytes per frame  :1070 */
    uint64_t data_offset;       /* dataChunkStart */
    uint64_t data_bytes;
} zl_wav_format;
/* Host-only chunk walk (no device needed); ZL_ERR_INVALID when WavAudioFormat::createReaderFor would return no reader
 * (:1718-1731) or the sample format is not one copySampleData decodes (:1232-1245). */
ZL_EXPORT int ZL_FN(wav_probe)(const void* wav, uint64_t nbytes, zl_wav_format* out);
/* ClipAudioSource_new(path) end to end (lib/libzl.cpp:107-128 -> lib/SamplerSynthSound.cpp:28-59): the file's bytes go to
 * the device as they are and are decoded there - left-justified int32, then f32(i) * 2^-31
 * (juce_AudioDataConverters.h:134,159,184,209; juce_AudioFormatReader.cpp:39-46,139-183) - into the clip's HBM frames;
 * the first min(2, channels) channels are kept; bytes the data chunk promises beyond the end of the image read as zero
 * (juce_WavAudioFormat.cpp:1212-1218).  No JUCE, no CPU decode. */
ZL_EXPORT int ZL_FN(clip_create_from_wav)(zl_ctx* ctx, const void* wav, uint64_t nbytes, int32_t* clip_id);
ZL_EXPORT int ZL_FN(clip_create_from_wav_file)(zl_ctx* ctx, const char* path, int32_t* clip_id);
/* The clip's sample data as SamplerSynthSound::audioData() holds it: planar[ch][length] floats, ch < min(2, channels). */
ZL_EXPORT int ZL_FN(clip_read_pcm)(zl_ctx* ctx, int32_t clip_id, float* const* planar, int32_t nch, int64_t length);
ZL_EXPORT int ZL_FN(clip_destroy)(zl_ctx* ctx, int32_t clip_id);
ZL_EXPORT int ZL_FN(clip_params_default)(zl_ctx* ctx, int32_t clip_id, zl_clip_params* out);
ZL_EXPORT int ZL_FN(clip_set_params)(zl_ctx* ctx, int32_t clip_id, const zl_clip_params* p);
ZL_EXPORT int ZL_FN(clip_get_params)(zl_ctx* ctx, int32_t clip_id, zl_clip_params* out);
/* (int)(getStartPosition(slice)*rate), (int)(getStopPosition(slice)*rate)
 * lib/SamplerSynthSound.cpp:96-104, lib/ClipAudioSource.cpp:261-277 */
ZL_EXPORT int ZL_FN(clip_slice_bounds)(zl_ctx* ctx, int32_t clip_id, int32_t slice,
                                       int32_t* start_pos, int32_t* stop_pos);

/* ---- commands: SamplerSynth::handleClipCommand ------------------------------------- */
/* Replaces SamplerSynth::handleClipCommand(cmd, currentTick) (lib/SamplerSynth.cpp:328-341): the
 * command is queued for bus midi_channel+2 of `session` and applied on the device, by the bus's own
 * warp, at the start of block `block_offset` of the next zl_render call
 * (SamplerChannel::process drain, lib/SamplerSynth.cpp:118-122; handleCommand :187-230). */
ZL_EXPORT int ZL_FN(command)(zl_ctx* ctx, int32_t session, int32_t block_offset,
                             uint64_t timestamp_tick, const zl_clip_command* cmd);
/* n commands at once (the per-period command stream of an offline render): arrays of length n. */
ZL_EXPORT int ZL_FN(command_batch)(zl_ctx* ctx, int32_t n, const int32_t* sessions, const int32_t* block_offsets,
                                   const uint64_t* timestamp_ticks, const zl_clip_command* cmds);
ZL_EXPORT int ZL_FN(clip_command_default)(zl_clip_command* cmd);
/* SamplerSynth::setChannelEnabled  lib/SamplerSynth.cpp:343-351 (bus = channel + 2). */
ZL_EXPORT int ZL_FN(bus_set_enabled)(zl_ctx* ctx, int32_t session, int32_t bus, int32_t enabled);
/* JackPassthrough_set{PanAmount,WetFx1Amount,WetFx2Amount,DryAmount,Muted}  lib/libzl.cpp:476-575 */
ZL_EXPORT int ZL_FN(passthrough_set)(zl_ctx* ctx, int32_t session, int32_t bus,
                                     const zl_passthrough_params* p);
ZL_EXPORT int ZL_FN(passthrough_get)(zl_ctx* ctx, int32_t session, int32_t bus,
                                     zl_passthrough_params* out);

/* ---- the hot path ------------------------------------------------------------------ */
/* One call = `nblocks` JACK periods of `nframes` frames for every bus of every session:
 * SamplerChannel::process (lib/SamplerSynth.cpp:116-148) -> SamplerSynthVoice::process
 * (lib/SamplerSynthVoice.cpp:174-270) or juce::SamplerVoice::renderNextBlock (J-mode)
 * -> JackPassthroughPrivate::process (lib/JackPassthrough.cpp:45-115)
 * -> per-block AudioLevels peak scan (lib/AudioLevels.cpp:357-385).
 * clocks: nblocks entries.  Asynchronous on the context stream. */
ZL_EXPORT int ZL_FN(render)(zl_ctx* ctx, int32_t nblocks, int32_t nframes,
                            const zl_block_clock* clocks);

/* Outputs of the last render (device -> caller's HOST buffers, n = nblocks*nframes floats each). */
ZL_EXPORT int ZL_FN(read_bus)(zl_ctx* ctx, int32_t session, int32_t bus, int32_t which,
                              float* left, float* right, int64_t n);
ZL_EXPORT int ZL_FN(read_master)(zl_ctx* ctx, int32_t session, float* left, float* right, int64_t n);
/* All sessions' masters at once: out is [n_sessions][2][n]. */
ZL_EXPORT int ZL_FN(read_master_all)(zl_ctx* ctx, float* out, int64_t n);
/* Pipelined variant for offline batch renders: enqueues the same copy on a second stream and returns; the next
 * zl_render computes into the other of two master buffers meanwhile.  out must be pinned (zl_host_alloc) and is
 * valid after zl_read_wait.  (The reference has no equivalent: JACK hands the period over synchronously.) */
ZL_EXPORT int ZL_FN(read_master_all_async)(zl_ctx* ctx, float* out, int64_t n);
/* The batched bounce (row n3 at scale): the same pipelined read-back, but every session's master arrives as 16- or 24-bit
 * PCM - the data chunk zl_bounce_wav(session, -1, ...) would write after its 104-byte header, i.e. the format of the
 * reference's recorder files (lib/AudioLevels.cpp:35-119 -> juce::WavAudioFormat, juce_AudioFormatWriter.cpp:64-79) -
 * encoded on the device: out is [n_sessions][n][L,R] little-endian samples, half (16) or three quarters (24) of the
 * bytes of the f32 planes.  out must be pinned; valid after zl_read_wait. */
ZL_EXPORT int ZL_FN(read_master_all_pcm_async)(zl_ctx* ctx, void* out, int64_t n, int32_t bits);
ZL_EXPORT int ZL_FN(read_wait)(zl_ctx* ctx);
/* Device pointers of the output planes, for zero-copy consumers (torch / NCCL):
 * bus planes are [n_buses_total][2][stride], master [n_sessions][2][stride]. */
ZL_EXPORT int ZL_FN(device_bus_ptr)(zl_ctx* ctx, int32_t which, float** d_ptr, int64_t* stride);
ZL_EXPORT int ZL_FN(device_master_ptr)(zl_ctx* ctx, float** d_ptr, int64_t* stride);
/* Per-block per-bus peaks of the last render: peaks[nblocks][n_buses_total][2] (int32),
 * sumsq[nblocks][n_buses_total][2] (double, may be NULL). */
ZL_EXPORT int ZL_FN(read_block_levels)(zl_ctx* ctx, int32_t* peaks, double* sumsq);
/* AudioLevels::timerCallback for one channel (lib/AudioLevels.cpp:347-412): decay by 10000, then
 * max with the peaks of block `block` of the last render (block < 0: no period captured). The
 * meter state lives in the context, one per bus. is_playback selects the hold logic (:394-398). */
ZL_EXPORT int ZL_FN(levels_tick)(zl_ctx* ctx, int32_t session, int32_t bus, int32_t block,
                                 int32_t is_playback, zl_levels* out);
/* Device voice table read-back: out[n_sessions*buses_per_session*voices_per_bus]. */
ZL_EXPORT int ZL_FN(read_voices)(zl_ctx* ctx, zl_voice_info* out, int32_t capacity);
/* ClipAudioSourcePositionsModel::peakGain raw max over the clip's playing voices
 * (lib/ClipAudioSourcePositionsModel.cpp:160-173, before hysteresis) + up to 32 (gain,progress). */
ZL_EXPORT int ZL_FN(read_positions)(zl_ctx* ctx, int32_t clip_id, float* gains, float* progress,
                                    int32_t capacity, int32_t* count);
/* ClipAudioSourcePositionsModel::peakGain (lib/ClipAudioSourcePositionsModel.cpp:160-173) as ClipAudioSource::syncAudioLevel
 * reads it (lib/ClipAudioSource.cpp:88-113): the maximum gain over the clip's live positions - the first 32 playing voices of
 * the clip in voice-table order (POSITION_COUNT, :5; createPositionID hands out slots in start order, :78-100, so the two
 * agree whenever at most 32 voices play one clip) - taken over only when it moved by more than 0.01 from the value last
 * reported (the model's hysteresis; the state lives in the context, one per clip).  first_progress: firstProgress() (:175-185),
 * -1 when nothing plays; may be NULL. */
ZL_EXPORT int ZL_FN(clip_peak_gain)(zl_ctx* ctx, int32_t clip_id, float* peak_gain, double* first_progress);
/* Source PCM floats spanned by the voices of the last render (unique samples incl. the +1 tap, summed per
 * voice per block; the algorithmic read traffic of SURVEY.md 8d is 4x this). */
ZL_EXPORT int ZL_FN(last_source_floats)(zl_ctx* ctx, uint64_t* n);
/* Running totals over the life of the context (voice-frames, source floats): lets a caller account a
 * sequence of asynchronous renders with one read. */
ZL_EXPORT int ZL_FN(total_counters)(zl_ctx* ctx, uint64_t* voice_frames, uint64_t* source_floats);
/* Number of voice-frames rendered by the last render (active voices x frames actually produced). */
ZL_EXPORT int ZL_FN(last_voice_frames)(zl_ctx* ctx, uint64_t* voice_frames);
/* Number of commands dropped (ring overflow lib/SamplerSynth.cpp:332-334 / unknown clip :330). */
ZL_EXPORT int ZL_FN(dropped_commands)(zl_ctx* ctx, uint64_t* n);
/* Optional device timing (bench.py roofline) with CUDA events on the context stream.  mode 1: every zl_render brackets
 * its whole plan/walk/gather launch sequence with one event pair; mode 2: one pair around each gather_kernel launch
 * (the dominant kernel); 0: off.  zl_kernel_time returns and clears the sum (ms) and the number of pairs. */
ZL_EXPORT int ZL_FN(kernel_timing)(zl_ctx* ctx, int32_t mode);
ZL_EXPORT int ZL_FN(kernel_time)(zl_ctx* ctx, double* total_ms, uint64_t* launches);
/* Host time spent inside zl_render since the last call (ms), the part of it blocked waiting for the device to release a
 * command slot (back-pressure, not host work), and the number of renders; clears the sums. */
ZL_EXPORT int ZL_FN(host_profile)(zl_ctx* ctx, double* submit_ms, double* wait_ms, uint64_t* renders);
/* Kernel launches issued by this context so far (bench `gpu_launches`). */
ZL_EXPORT int ZL_FN(launch_count)(zl_ctx* ctx, uint64_t* n);

/* ---- WaveFormItem / juce::AudioThumbnail ------------------------------------------- */
/* Level 0: LevelDataSource::readNextBlock -> readMaxLevels -> MinMaxValue::setFloat
 * (JUCE/juce_audio_utils/gui/juce_AudioThumbnail.cpp:225-269, :46-66; sizing :666-670).
 * Builds nch x (1 + len/samples_per_thumb) int8 (min,max) pairs on the device. */
ZL_EXPORT int ZL_FN(thumb_build)(zl_ctx* ctx, int32_t clip_id, int32_t samples_per_thumb);
ZL_EXPORT int ZL_FN(thumb_size)(zl_ctx* ctx, int32_t clip_id, int32_t* n_thumb, int32_t* nch);
/* out: [nch][n_thumb][2] int8 */
ZL_EXPORT int ZL_FN(thumb_read_level0)(zl_ctx* ctx, int32_t clip_id, int8_t* out);
/* n4: the thumbnail in JUCE's `jatm` stream format - AudioThumbnail::saveTo / loadFrom
 * (JUCE/juce_audio_utils/gui/juce_AudioThumbnail.cpp:593-639), what AudioThumbnailCache::storeThumb keeps per file hash:
 * "jatm", samplesPerThumbSample i32, totalSamples i64, numSamplesFinished i64 (= whole thumb blocks read, :735-743),
 * numThumbnailSamples i32, numChannels i32, sampleRate i32, 16 reserved bytes, then [thumb sample][channel](min, max) int8.
 * save: the device level-0 table is transposed on the device into that order (builds it at 512 samples per thumb
 * sample if the clip has none, lib/WaveFormItem.cpp:22); out == NULL or cap too small: *size set, ZL_ERR_STATE.
 * load: replaces the clip's level-0 table (no pass over the PCM); the stream must describe this clip (channels and
 * 1 + length / samplesPerThumbSample entries), anything else is ZL_ERR_INVALID. */
ZL_EXPORT int ZL_FN(thumb_save_jatm)(zl_ctx* ctx, int32_t clip_id, void* out, uint64_t cap, uint64_t* size);
ZL_EXPORT int ZL_FN(thumb_load_jatm)(zl_ctx* ctx, int32_t clip_id, const void* jatm, uint64_t nbytes);
/* CachedWindow::refillCache (juce_AudioThumbnail.cpp:430-521) for all channels:
 * out [nch][width][2] int8 (min,max).  Returns ZL_ERR_INVALID where the reference returns false. */
ZL_EXPORT int ZL_FN(thumb_pixels)(zl_ctx* ctx, int32_t clip_id, double t0, double t1,
                                  int32_t width, int8_t* out);
/* CachedWindow::drawChannel geometry + QPainterContext::fillRectList integer truncation
 * (juce_AudioThumbnail.cpp:379-422, lib/QPainterContext.cpp:190-200): headless mask[height][width]
 * (1 = brush colour) for `channel`, area = (0,0,width,height), vertical zoom `zoom`. */
ZL_EXPORT int ZL_FN(thumb_raster)(zl_ctx* ctx, int32_t clip_id, int32_t channel, double t0,
                                  double t1, int32_t width, int32_t height, float zoom,
                                  uint8_t* mask);

/* ---- standalone stages over caller-owned buffers (K4/K5 as separate operators) ------- */
/* JackPassthroughPrivate::process on HOST buffers in[2][n] -> out[6][n]
 * (dryL,dryR,fx1L,fx1R,fx2L,fx2R). */
ZL_EXPORT int ZL_FN(passthrough_process)(zl_ctx* ctx, const zl_passthrough_params* p,
                                         const float* in_l, const float* in_r, int64_t n,
                                         float* out6);
/* Peak scan of one period (lib/AudioLevels.cpp:365-382) + sum of squares, HOST buffers. */
ZL_EXPORT int ZL_FN(levels_scan)(zl_ctx* ctx, const float* l, const float* r, int64_t n,
                                 int32_t* peak_a, int32_t* peak_b, double* sumsq_a, double* sumsq_b);
/* The same scan without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that may not call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Input staging and accumulators live in a file-scope static __device__ array of 256 + 8 * ZL_STATIC_SCAN_MAX_FRAMES bytes
 * (2 MiB + 256 B), held by every GPU that loads the library until the process exits; the call makes no cudaMalloc, cudaFree
 * or allocating library call.  n > ZL_STATIC_SCAN_MAX_FRAMES is refused with ZL_ERR_INVALID (zl_last_error says why).
 * All calls on a device share the one array (the accumulators are cleared at every call): calls that overlap on the GPU -
 * two contexts or two streams scanning at the same time on one device - race; serialise them.  Same results as
 * zl_levels_scan: peaks bit-equal, sums of squares equal up to f64 summation order. */
#define ZL_STATIC_SCAN_MAX_FRAMES 262144
ZL_EXPORT int ZL_FN(levels_scan_static)(zl_ctx* ctx, const float* l, const float* r, int64_t n,
                                        int32_t* peak_a, int32_t* peak_b, double* sumsq_a, double* sumsq_b);

/* ---- n3: bounce ------------------------------------------------------ */
/* The last render of one stereo port as the WAV file the reference's DiskWriter leaves on disk (lib/AudioLevels.cpp:35-119:
 * WavAudioFormat::createWriterFor(stream, rate, 2 channels, 16 bit), AudioFormatWriter::writeFromFloatArrays): floats ->
 * int32 by convertFloatsToInts (juce_AudioFormatWriter.cpp:64-79), top `bits` (16 or 24) kept
 * (juce_AudioDataConverters.h:159,184), behind the writer's RIFF | JUNK | fmt | data header (juce_WavAudioFormat.cpp:1391-1500);
 * converted on the device, one D2H copy of the packed samples.  bus < 0: the session's master mix (ZL_OUT_MASTER), else
 * plane `which` (0 raw, 1 dry, 2 fx1, 3 fx2) of that bus.  out == NULL or cap too small: *size is set, ZL_ERR_STATE. */
/* The same writer over HOST floats (r == NULL: a mono file): upload, convert on the device, header on the host. */
ZL_EXPORT int ZL_FN(wav_encode)(zl_ctx* ctx, const float* l, const float* r, int64_t n, double rate, int32_t bits,
                                void* out, uint64_t cap, uint64_t* size);
ZL_EXPORT int ZL_FN(bounce_wav)(zl_ctx* ctx, int32_t session, int32_t bus, int32_t which, int32_t bits,
                                void* out, uint64_t cap, uint64_t* size);

/* ---- n2: the command-scheduling front-end ----------------------- */
/* SyncTimer's 96-ppqn step ring (32768 steps) as a host object: commands are scheduled `delay` steps ahead
 * (SyncTimer::scheduleClipCommand, lib/SyncTimer.cpp:1011-1048, same-step merge of equivalentTo commands), and
 * zl_sched_advance walks the steps that fall due in each of `nblocks` JACK periods the way SyncTimerPrivate::process does
 * (:452-702): it fills playhead / playhead_usecs / subbeat_usecs of clocks[] (the caller sets current_usecs / next_usecs, what
 * jack_get_cycle_times would report) and returns the commands that fire, each with its period and the tick
 * SamplerSynth::handleClipCommand is given (:553-558) - ready for zl_command_batch + zl_render.  Pure host logic. */
typedef struct zl_sched zl_sched;
enum { ZL_SCHED_CLIP_COMMAND = 0, ZL_SCHED_CHANNEL_ENABLED = 1 };
typedef struct {
    int32_t kind;               /* ZL_SCHED_* */
    int32_t block;              /* period of this advance the event belongs to */
    uint64_t tick;              /* jackPlayhead when it fired */
    zl_clip_command cmd;        /* ZL_SCHED_CLIP_COMMAND */
    int32_t channel, enabled;   /* ZL_SCHED_CHANNEL_ENABLED: SamplerSynth::setChannelEnabled(channel, enabled) */
    int32_t reserved;
} zl_sched_event;
ZL_EXPORT int ZL_FN(sched_create)(double sample_rate, int32_t buffer_size, zl_sched** out);
ZL_EXPORT int ZL_FN(sched_destroy)(zl_sched* s);
ZL_EXPORT int ZL_FN(sched_start)(zl_sched* s, int32_t bpm);                       /* SyncTimer::start  :870-879 */
ZL_EXPORT int ZL_FN(sched_stop)(zl_sched* s);                                     /* SyncTimer::stop   :881-924: what still waits is muted and re-queued one step behind the read head */
ZL_EXPORT int ZL_FN(sched_set_bpm)(zl_sched* s, int32_t bpm);                     /* SyncTimer::setBpm :955-975 */
ZL_EXPORT int ZL_FN(sched_schedule)(zl_sched* s, const zl_clip_command* cmd, uint64_t delay);
ZL_EXPORT int ZL_FN(sched_schedule_channel_enabled)(zl_sched* s, uint64_t delay, int32_t channel, int32_t enabled);
ZL_EXPORT int ZL_FN(sched_queue_start)(zl_sched* s, int32_t clip, int32_t midi_channel);   /* queueClipToStartOnChannel :816-832 */
ZL_EXPORT int ZL_FN(sched_queue_stop)(zl_sched* s, int32_t clip, int32_t midi_channel);    /* queueClipToStopOnChannel  :834-860 */
/* *count = events produced (also when cap was too small: ZL_ERR_STATE, and the ring has moved on - size cap generously) */
ZL_EXPORT int ZL_FN(sched_advance)(zl_sched* s, int32_t nblocks, int32_t nframes, zl_block_clock* clocks,
                                   zl_sched_event* events, int32_t cap, int32_t* count);
ZL_EXPORT int ZL_FN(sched_state)(zl_sched* s, uint64_t* playhead, uint64_t* playhead_usecs, uint64_t* subbeat_usecs,
                                 uint64_t* cumulative_beat, int32_t* beat, uint64_t* schedule_ahead);

/* ---- one mix spanning GPUs ------------------------------------- */
/* The only exchange step on the path: when the dry ports of buses rendered on different GPUs feed ONE stereo pair (what JACK
 * does when every track's dryOut is connected to the same system:playback ports, lib/JackPassthrough.cpp:94-101), the
 * per-rank master planes are summed over NCCL (NVLink / NVSwitch).  One process per GPU; the NCCL C API is taken from the
 * libnccl.so.2 already in the process or on the loader path (ZL_B200_NCCL_LIB overrides), never linked at build time.
 * zl_comm_unique_id: rank 0 of a mix group fills 128 bytes (ncclUniqueId) and hands them to the others out of band.
 * zl_comm_init: joins the group (collective: every rank of the group calls it).
 * zl_mix_reduce: float32 sum of the master planes [n_sessions][2][stride] of the LAST render, in place on the context stream
 *   (so zl_read_master* / zl_bounce_wav that follow see the spanned mix); root >= 0: ncclReduce, only that rank holds the sum
 *   (the others keep their own partial mix); root < 0: ncclAllReduce.  The summation order differs from the serial
 *   reference: results agree to float reassociation (1e-5 of full scale), not bit for bit; per-bus planes and meters are
 *   untouched (a meter of the spanned mix must be taken after the reduce: zl_levels_scan). */
ZL_EXPORT int ZL_FN(comm_unique_id)(void* id128);
ZL_EXPORT int ZL_FN(comm_init)(zl_ctx* ctx, const void* nccl_unique_id, int32_t rank, int32_t nranks);
ZL_EXPORT int ZL_FN(mix_reduce)(zl_ctx* ctx, int32_t root);
ZL_EXPORT int ZL_FN(comm_destroy)(zl_ctx* ctx);

/* ---- pinned host memory ------------------------------------------------------------- */
/* Page-locked host buffers for callers that want full-rate PCIe transfers on zl_read_* / clip upload
 * (the JACK port buffers of the reference, lib/SamplerSynth.cpp:132-133, are plain shared memory). */
ZL_EXPORT int ZL_FN(host_alloc)(void** p, size_t bytes);
ZL_EXPORT int ZL_FN(host_free)(void* p);

#ifdef __cplusplus
}
#endif
#endif /* ZL_B200_H */
