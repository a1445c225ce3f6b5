/*
 * include/zl/libzl_facade.h — the reference's C++ classes for the sampler render path, kept by name and method so that
 * host code written against libzl compiles against this library: SamplerSynth, SamplerSynthVoice, ClipAudioSource,
 * ClipCommand, JackPassthrough, AudioLevels, WaveFormItem, ClipAudioSourcePositionsModel, SyncTimer.
 *
 * They are thin host objects over the C ABI of include/zl_b200.h (the CUDA library); no Qt, no JACK, no tracktion:
 *   QObject -> plain class, QString -> std::string, QVariantList -> std::vector<double>, jack_default_audio_sample_t -> float,
 *   quint64 -> uint64_t, te::Engine* -> ignored, Q_SIGNAL -> std::function callbacks where the reference exposes one.
 * Where the reference is driven by JACK's process callback, the host calls SamplerSynth::process (one period) or
 * SamplerSynth::render (several periods, the offline form) itself.
 *
 * Each class cites the reference header it mirrors (paths relative to /root/reference).
 */
#ifndef ZL_LIBZL_FACADE_H
#define ZL_LIBZL_FACADE_H

#include <cstdint>
#include <functional>
#include <memory>
#include <string>
#include <vector>

#include "../zl_b200.h"

#define ZL_FACADE_API __attribute__((visibility("default")))

class ClipAudioSource;
class SamplerSynth;
class SyncTimer;
namespace tracktion_engine { class Engine; }

/* lib/ClipCommand.h:11-107 */
struct ZL_FACADE_API ClipCommand {
    ClipCommand() {}
    ClipCommand(ClipAudioSource* clip, int midiNote) : clip(clip), midiNote(midiNote) {}
    ClipAudioSource* clip{nullptr};
    int midiNote{-1};
    int midiChannel{-1};
    bool startPlayback{false};
    bool stopPlayback{false};
    bool changeSlice{false};
    int slice{-1};
    bool changeLooping{false};
    bool looping{false};
    bool changePitch{false};
    float pitchChange{0.0f};
    bool changeSpeed{false};
    float speedRatio{0.0f};
    bool changeGainDb{false};
    float gainDb{0.0f};
    bool changeVolume{false};
    float volume{0.0f};
    bool equivalentTo(ClipCommand* other) const;
    static ClipCommand* noEffectCommand(ClipAudioSource* clip);
    static ClipCommand* effectedCommand(ClipAudioSource* clip);
    static ClipCommand* channelCommand(ClipAudioSource* clip, int channelID);
    static void clear(ClipCommand* command);
};

/* juce::ADSR::Parameters  JUCE/juce_audio_basics/utilities/juce_ADSR.h:52-74 (defaults 0.1, 0.1, 1.0, 0.1) */
struct ADSRParameters {
    ADSRParameters() = default;
    ADSRParameters(float a, float d, float s, float r) : attack(a), decay(d), sustain(s), release(r) {}
    float attack = 0.1f, decay = 0.1f, sustain = 1.0f, release = 0.1f;
};

/* lib/ClipAudioSourcePositionsModel.h:9-52 - the view ClipAudioSource::playbackPositionsModel() hands out */
class ZL_FACADE_API ClipAudioSourcePositionsModel {
public:
    explicit ClipAudioSourcePositionsModel(ClipAudioSource* clip) : m_clip(clip) {}
    float peakGain() const;                 /* max over the live positions, 0.01 hysteresis  .cpp:160-173 */
    double firstProgress() const;           /* -1 when nothing plays                         .cpp:175-185 */
    int rowCount() const;                   /* live positions (at most 32) */
    void positions(std::vector<float>& gains, std::vector<float>& progress) const;
private:
    ClipAudioSource* m_clip;
};

/* lib/ClipAudioSource.h:28-237 */
class ZL_FACADE_API ClipAudioSource {
public:
    /* filepath: a WAV file (the library's own reader: juce_WavAudioFormat's rules); the clip registers itself with
     * SamplerSynth::instance() like the reference's constructor does (lib/ClipAudioSource.cpp:199). */
    explicit ClipAudioSource(tracktion_engine::Engine* engine, SyncTimer* syncTimer, const char* filepath, bool muted = false);
    /* in-memory variant: planar float PCM */
    ClipAudioSource(const float* const* planar, int channels, int64_t frames, double sampleRate, bool muted = false);
    ~ClipAudioSource();

    void setProgressCallback(void (*functionPtr)(float));
    void syncProgress();
    void setStartPosition(float startPositionInSeconds);
    float getStartPosition(int slice = -1) const;
    float getStopPosition(int slice = -1) const;
    void setLooping(bool looping);
    bool getLooping() const;
    void setLength(float beat, int bpm);
    float getLengthInBeats() const;
    void setPitch(float pitchChange, bool immediate = false);       /* stored; the tracktion proxy render is out of scope */
    void setSpeedRatio(float speedRatio, bool immediate = false);   /* stored */
    void setGain(float db);                                         /* stored */
    void setVolume(float vol);
    void setVolumeAbsolute(float vol);
    float volumeAbsolute() const;
    void setAudioLevelChangedCallback(void (*functionPtr)(float));
    void play(bool loop = true, int midiChannel = -2);
    void stop(int midiChannel = -3);
    float getDuration();
    const char* getFileName() const;
    const char* getFilePath() const;

    int id() const;
    void setId(int id);
    float audioLevel() const;
    double position() const;
    ClipAudioSourcePositionsModel* playbackPositionsModel();

    int slices() const;
    void setSlices(int slices);
    std::vector<double> slicePositions() const;
    void setSlicePositions(const std::vector<double>& slicePositions);
    double slicePosition(int slice) const;
    void setSlicePosition(int slice, float position);
    int sliceBaseMidiNote() const;
    void setSliceBaseMidiNote(int sliceBaseMidiNote);
    int sliceForMidiNote(int midiNote) const;
    int keyZoneStart() const;
    void setKeyZoneStart(int keyZoneStart);
    int keyZoneEnd() const;
    void setKeyZoneEnd(int keyZoneEnd);
    int rootNote() const;
    void setRootNote(int rootNote);
    float pan();
    void setPan(float pan);

    float adsrAttack() const;
    void setADSRAttack(const float& newValue);      /* like the reference: the other three fall back to JUCE's defaults */
    float adsrDecay() const;
    void setADSRDecay(const float& newValue);
    float adsrSustain() const;
    void setADSRSustain(const float& newValue);
    float adsrRelease() const;
    void setADSRRelease(const float& newValue);
    const ADSRParameters& adsrParameters() const;
    void setADSRParameters(const ADSRParameters& parameters);

    /* ClipAudioSource::Private::syncAudioLevel (lib/ClipAudioSource.cpp:88-113): call at UI rate */
    void syncAudioLevel();
    int32_t deviceClip() const;                    /* the zl_b200 clip id behind this object */
private:
    class Private;
    Private* d;
    friend class SamplerSynth;
    friend class ClipAudioSourcePositionsModel;
    ClipAudioSource(const ClipAudioSource&) = delete;
    ClipAudioSource& operator=(const ClipAudioSource&) = delete;
};

/* lib/SamplerSynthVoice.h:9-34: a view of one slot of the device-resident voice table */
class ZL_FACADE_API SamplerSynthVoice {
public:
    SamplerSynthVoice(SamplerSynth* synth, int channelIndex, int slot) : m_synth(synth), m_channel(channelIndex), m_slot(slot) {}
    void refresh();                                 /* reads the slot back (zl_read_voices); the getters below use it */
    bool isPlaying{false};
    double sourceSamplePosition() const { return m_info.position; }
    double pitchRatio() const { return m_info.pitch_ratio; }
    float envelope() const { return m_info.envelope; }
    int midiNote() const { return m_info.midi_note; }
    int32_t deviceClip() const { return m_info.clip; }
private:
    SamplerSynth* m_synth; int m_channel, m_slot;
    zl_voice_info m_info{};
};

/* lib/SamplerSynth.h:14-57 */
class ZL_FACADE_API SamplerSynth {
public:
    static SamplerSynth* instance();
    explicit SamplerSynth();
    ~SamplerSynth();

    /* allocates the device tables: 12 SamplerChannels x 8 voices (lib/SamplerSynth.cpp:254-278).  sampleRate replaces
     * jack_get_sample_rate; maxFrames / maxBlocks bound a render call.  Returns false (and keeps lastError()) without a
     * CUDA device: there is no CPU fallback. */
    bool initialize(tracktion_engine::Engine* engine, double sampleRate = 48000.0, int maxFrames = 1024, int maxBlocks = 64);
    tracktion_engine::Engine* engine() const;
    bool isInitialized() const;
    const char* lastError() const;

    void registerClip(ClipAudioSource* clip);
    void unregisterClip(ClipAudioSource* clip);

    /* takes ownership of the command (the reference returns it to SyncTimer's pool) */
    void handleClipCommand(ClipCommand* clipCommand);
    void handleClipCommand(ClipCommand* clipCommand, uint64_t currentTick);
    /* same, applied at the start of period `block` of the next render (offline scheduling) */
    void handleClipCommandAt(ClipCommand* clipCommand, uint64_t currentTick, int block);
    void setChannelEnabled(const int& channel, const bool& enabled) const;

    /* fraction of the rendered audio time the device needed for the last process()/render(): what jack_cpu_load reports
     * for the reference's first channel (lib/SamplerSynth.cpp:143-145, 320-326), as a percentage */
    float cpuLoad() const;

    /* the JACK cycle: every channel's SamplerChannel::process + its JackPassthrough + the meters' scan */
    bool process(uint32_t nframes, const zl_block_clock& clock);
    bool render(int nblocks, uint32_t nframes, const zl_block_clock* clocks);
    /* ports of the last cycle: which = ZL_BUS_RAW (left_out/right_out), ZL_BUS_DRY, ZL_BUS_FX1, ZL_BUS_FX2 */
    bool readChannel(int channel, int which, float* left, float* right) const;
    bool readMaster(float* left, float* right) const;
    int64_t framesInLastRender() const;
    int blocksInLastRender() const;
    uint64_t renderSerial() const;                  /* counts process()/render() calls */

    SamplerSynthVoice voice(int channel, int slot) { return SamplerSynthVoice(this, channel + 2, slot); }
    zl_ctx* context() const;
private:
    class Private;
    Private* d{nullptr};
    friend class SamplerSynthVoice;
};

/* lib/SyncTimer.h:19-223 - the clip-command half of the timer (row n2): the 96-ppqn step ring, scheduleClipCommand with its
 * same-step merge, clips queued to the next bar, start/stop/bpm, the playhead getters the voices read, and the
 * SamplerChannelEnabledStateOperation timer command.  The MIDI-note half (scheduleNote, scheduleMidiBuffer, send*Immediately),
 * the JACK transport and the hardware-device signals are outside SURVEY.md section 8; stopClip is declared by the
 * reference but never defined (lib/SyncTimer.cpp has no body for it) and is left out.
 * Where the reference is driven by JACK (SyncTimerPrivate::process, lib/SyncTimer.cpp:452-702) and by its timer thread,
 * the host calls process()/render(): the steps that fall due are walked, their commands go to
 * SamplerSynth::instance() with their period and tick, channel switches are applied, and the periods are rendered. */
class ZL_FACADE_API SyncTimer {
public:
    static SyncTimer* instance();
    explicit SyncTimer(double sampleRate = 48000.0, int bufferSize = 1024);
    ~SyncTimer();
    void queueClipToStart(ClipAudioSource* clip);
    void queueClipToStop(ClipAudioSource* clip);
    void queueClipToStartOnChannel(ClipAudioSource* clip, int midiChannel);
    void queueClipToStopOnChannel(ClipAudioSource* clip, int midiChannel);
    void start(int bpm);
    void stop();
    int getInterval(int bpm);
    float subbeatCountToSeconds(uint64_t bpm, uint64_t beats) const;
    uint64_t secondsToSubbeatCount(uint64_t bpm, float seconds) const;
    int getMultiplier();
    uint64_t getBpm() const;
    void setBpm(uint64_t bpm);
    uint64_t scheduleAheadAmount() const;
    int beat() const;
    uint64_t cumulativeBeat() const;
    const uint64_t& jackPlayhead() const;
    const uint64_t& jackPlayheadUsecs() const;
    const uint64_t& jackSubbeatLengthInMicroseconds() const;
    /* takes ownership of the command (merged into an equivalent one waiting on the same step, or queued) */
    void scheduleClipCommand(ClipCommand* command, uint64_t delay);
    /* operation: TimerCommand::SamplerChannelEnabledStateOperation (8) - parameter1 the channel (-2..9), parameter2 0 = off.
     * Other operations are not part of the path and are ignored. */
    void scheduleTimerCommand(uint64_t delay, int operation, int parameter1 = 0, int parameter2 = 0, int parameter3 = 0);
    bool timerRunning();
    ClipCommand* getClipCommand();
    void deleteClipCommand(ClipCommand* command);
    /* Q_SIGNAL clipCommandSent(ClipCommand*): called for every command handed to the sampler (the pointer is valid for the call) */
    std::function<void(ClipCommand*)> clipCommandSent;

    /* one JACK cycle (current_usecs / next_usecs as jack_get_cycle_times gives them), or `nblocks` cycles back to back
     * starting at firstUsecs with periods of nframes * 1e6 / sampleRate microseconds - the offline form: the scheduler runs
     * ahead over all of them and the device renders them as one batch (split where a channel is switched). */
    bool process(uint32_t nframes, uint64_t currentUsecs, uint64_t nextUsecs);
    bool render(int nblocks, uint32_t nframes, uint64_t firstUsecs);
    /* commands handed over / periods walked so far */
    uint64_t commandsSent() const;
    uint64_t periodsProcessed() const;
    zl_sched* scheduler() const;
private:
    class Private;
    Private* d{nullptr};
};

/* lib/JackPassthrough.h:26-58: the passthrough of one SamplerSynth channel (-2, -1 global; 0..9 tracks) */
class ZL_FACADE_API JackPassthrough {
public:
    explicit JackPassthrough(int channel);
    float dryAmount() const;
    void setDryAmount(const float& newValue);
    float wetFx1Amount() const;
    void setWetFx1Amount(const float& newValue);
    float wetFx2Amount() const;
    void setWetFx2Amount(const float& newValue);
    float panAmount() const;
    void setPanAmount(const float& newValue);
    bool muted() const;
    void setMuted(const bool& newValue);
private:
    int m_bus;
    zl_passthrough_params get() const;
    void set(const zl_passthrough_params& p);
};

/* lib/AudioLevels.h:39-227 */
class ZL_FACADE_API AudioLevels {
public:
    static AudioLevels* instance();
    float add(float db1, float db2);
    /* AudioLevels::timerCallback (lib/AudioLevels.cpp:347-412) on the last rendered period: the ten track meters
     * (channels 0..9), playback = the global un-effected channel.  Call at the reference's 20 Hz. */
    void timerCallback();
    float captureA{-200.0f}, captureB{-200.0f};
    float playbackA{-200.0f}, playbackB{-200.0f}, playbackAHold{-200.0f}, playbackBHold{-200.0f}, playback{-200.0f};
    float recordingA{-200.0f}, recordingB{-200.0f};
    float channelsA[10] = {0}, channelsB[10] = {0};
    std::vector<float> channels() const;            /* getChannelsAudioLevels: add(channelsA[i], channelsB[i]) */
    /* recorder (lib/AudioLevels.cpp:35-119, 514-569): the last render of a channel's dry port as the 16-bit WAV file
     * DiskWriter leaves */
    void setChannelToRecord(int channel, bool shouldRecord = true);
    void setChannelFilenamePrefix(int channel, const std::string& fileNamePrefix);
    void setRecordGlobalPlayback(bool shouldRecord = true);
    void setGlobalPlaybackFilenamePrefix(const std::string& fileNamePrefix);
    void startRecording();                          /* writes the files for the last render and returns */
    void stopRecording() {}
    bool isRecording() const { return false; }
private:
    AudioLevels() {}
    std::vector<float> m_levels = std::vector<float>(10, 0.0f);
    bool m_recordChannel[10] = {false}; std::string m_prefix[10];
    bool m_recordGlobal = false; std::string m_globalPrefix;
    uint64_t m_scannedSerial = 0;
};

/* lib/WaveFormItem.h:17-71: the item's data side; paint() becomes a headless mask (Qt itself: unavailable on this box) */
class ZL_FACADE_API WaveFormItem {
public:
    WaveFormItem() {}
    std::string source() const { return m_source; }
    void setSource(const std::string& source);      /* a WAV path */
    void setSource(ClipAudioSource* clip);          /* or a clip that is already resident */
    double length() const;
    uint32_t color() const { return m_color; }
    void setColor(uint32_t argb) { m_color = argb; }
    double start() const { return m_start; }
    void setStart(double start) { m_start = start; }
    double end() const { return m_end; }
    void setEnd(double end) { m_end = end; }
    /* WaveFormItem::paint (lib/WaveFormItem.cpp:130-143): drawChannel(0, start, min(end, length), zoom 1) into
     * mask[height][width] (1 = brush colour); false where the reference draws nothing */
    bool paint(int width, int height, std::vector<uint8_t>& mask) const;
    /* the same as packed ARGB pixels (background 0) */
    bool paintARGB(int width, int height, std::vector<uint32_t>& pixels) const;
private:
    std::string m_source;
    std::unique_ptr<ClipAudioSource> m_owned;
    ClipAudioSource* m_clip = nullptr;
    uint32_t m_color = 0xffffffffu;
    double m_start = 0, m_end = 0;
};

#endif /* ZL_LIBZL_FACADE_H */
