// libzl_b200/host/libzl_facade.cpp — the reference's C++ classes (include/zl/libzl_facade.h) forwarding to the C ABI of
// include/zl_b200.h.  Host-side glue only: every sample is rendered by the CUDA library; without a device
// SamplerSynth::initialize fails (no CPU fallback).  Reference lines are cited where behaviour is mirrored.
#include "../../include/zl/libzl_facade.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>

namespace {
float decibelsToVolumeFaderPosition(float db) {    // tracktion_engine/utilities/tracktion_AudioUtilities.cpp:138-141
    return db > -100.0f ? std::exp((db - 6.0f) * (1.0f / 20.0f)) : 0.0f;
}
double gainToDecibels(double gain) { return gain > 0.0 ? std::max(-100.0, std::log10(gain) * 20.0) : -100.0; }   // juce::Decibels
double decibelsToGain(double db) { return db > -100.0 ? std::pow(10.0, db * 0.05) : 0.0; }
}  // namespace

// ---- ClipCommand  lib/ClipCommand.h:33-106 -------------------------------------------------------------------------
bool ClipCommand::equivalentTo(ClipCommand* other) const {
    return clip == other->clip
        && ((changeSlice == true && other->changeSlice == true && slice == other->slice)
            || (changeSlice == false && other->changeSlice == false && midiNote == other->midiNote && midiChannel == other->midiChannel));
}
ClipCommand* ClipCommand::noEffectCommand(ClipAudioSource* clip) { auto* c = new ClipCommand; c->clip = clip; c->midiChannel = -2; c->midiNote = 60; return c; }
ClipCommand* ClipCommand::effectedCommand(ClipAudioSource* clip) { auto* c = new ClipCommand; c->clip = clip; c->midiChannel = -1; c->midiNote = 60; return c; }
ClipCommand* ClipCommand::channelCommand(ClipAudioSource* clip, int channelID) { auto* c = new ClipCommand; c->clip = clip; c->midiChannel = channelID; return c; }
void ClipCommand::clear(ClipCommand* command) { ClipAudioSource* keep = nullptr; *command = ClipCommand(); command->clip = keep; }

// ---- SamplerSynth ------------------------------------------------------------------------------------------------
class SamplerSynth::Private {
public:
    zl_ctx* ctx = nullptr;
    tracktion_engine::Engine* engine = nullptr;
    double rate = 48000.0;
    std::string err;
    int nblocks = 0, nframes = 0;
    uint64_t serial = 0;
    float cpuLoad = 0.0f;
};

SamplerSynth* SamplerSynth::instance() {
    static SamplerSynth* instance{nullptr};
    if (!instance) instance = new SamplerSynth();
    return instance;
}
SamplerSynth::SamplerSynth() : d(new Private) {}
SamplerSynth::~SamplerSynth() { if (d->ctx) zl_ctx_destroy(d->ctx); delete d; }

bool SamplerSynth::initialize(tracktion_engine::Engine* engine, double sampleRate, int maxFrames, int maxBlocks) {
    d->engine = engine;
    if (d->ctx) return true;
    zl_config cfg;
    zl_config_default(&cfg);                        // 12 channels x 8 voices, Z-mode, quirks faithful  lib/SamplerSynth.cpp:23,258
    cfg.sample_rate = sampleRate; cfg.max_frames = maxFrames; cfg.max_blocks = maxBlocks; cfg.outputs = ZL_OUT_ALL;
    const int rc = zl_ctx_create(&cfg, &d->ctx);
    if (rc != ZL_OK) {
        d->ctx = nullptr;
        d->err = rc == ZL_ERR_NODEVICE ? "no CUDA device: libzl_b200 has no CPU fallback" : "zl_ctx_create failed";
        return false;
    }
    d->rate = sampleRate;
    zl_kernel_timing(d->ctx, 1);                    // cpuLoad(): device time of each render
    return true;
}
tracktion_engine::Engine* SamplerSynth::engine() const { return d->engine; }
bool SamplerSynth::isInitialized() const { return d->ctx != nullptr; }
const char* SamplerSynth::lastError() const { return d->ctx ? zl_last_error(d->ctx) : d->err.c_str(); }
zl_ctx* SamplerSynth::context() const { return d->ctx; }

void SamplerSynth::registerClip(ClipAudioSource*) {}        // the clip table entry is made by the clip's constructor (it owns the PCM upload)
void SamplerSynth::unregisterClip(ClipAudioSource* clip) { if (d->ctx && clip && clip->deviceClip() >= 0) zl_clip_destroy(d->ctx, clip->deviceClip()); }

void SamplerSynth::handleClipCommand(ClipCommand* clipCommand) { handleClipCommand(clipCommand, 0); }
void SamplerSynth::handleClipCommand(ClipCommand* clipCommand, uint64_t currentTick) { handleClipCommandAt(clipCommand, currentTick, 0); }
void SamplerSynth::handleClipCommandAt(ClipCommand* c, uint64_t currentTick, int block) {
    if (!c) return;
    if (d->ctx && c->clip) {                        // lib/SamplerSynth.cpp:328-341 (unknown clip / channel: dropped by zl_command)
        zl_clip_command k;
        zl_clip_command_default(&k);
        k.clip = c->clip->deviceClip(); k.midi_note = c->midiNote; k.midi_channel = c->midiChannel;
        k.start_playback = c->startPlayback; k.stop_playback = c->stopPlayback; k.change_slice = c->changeSlice; k.slice = c->slice;
        k.change_looping = c->changeLooping; k.looping = c->looping; k.change_pitch = c->changePitch; k.pitch_change = c->pitchChange;
        k.change_speed = c->changeSpeed; k.speed_ratio = c->speedRatio; k.change_gain_db = c->changeGainDb; k.gain_db = c->gainDb;
        k.change_volume = c->changeVolume; k.volume = c->volume;
        zl_command(d->ctx, 0, block, currentTick, &k);
    }
    delete c;                                       // ownership was handed over (the reference returns it to SyncTimer's pool)
}
void SamplerSynth::setChannelEnabled(const int& channel, const bool& enabled) const {
    if (d->ctx && channel > -3 && channel < 10) zl_bus_set_enabled(d->ctx, 0, channel + 2, enabled ? 1 : 0);   // lib/SamplerSynth.cpp:343-351
}
float SamplerSynth::cpuLoad() const { return d->cpuLoad; }

bool SamplerSynth::process(uint32_t nframes, const zl_block_clock& clock) { return render(1, nframes, &clock); }
bool SamplerSynth::render(int nblocks, uint32_t nframes, const zl_block_clock* clocks) {
    if (!d->ctx || zl_render(d->ctx, nblocks, (int32_t) nframes, clocks) != ZL_OK) return false;
    d->nblocks = nblocks; d->nframes = (int) nframes; d->serial++;
    double ms = 0.0; uint64_t n = 0;
    if (zl_kernel_time(d->ctx, &ms, &n) == ZL_OK && n)      // waits for the render: a JACK cycle hands its buffers over synchronously too
        d->cpuLoad = (float) (100.0 * ms / (1000.0 * (double) nblocks * nframes / d->rate));
    return true;
}
int64_t SamplerSynth::framesInLastRender() const { return (int64_t) d->nblocks * d->nframes; }
int SamplerSynth::blocksInLastRender() const { return d->nblocks; }
uint64_t SamplerSynth::renderSerial() const { return d->serial; }
bool SamplerSynth::readChannel(int channel, int which, float* left, float* right) const {
    return d->ctx && zl_read_bus(d->ctx, 0, channel + 2, which, left, right, framesInLastRender()) == ZL_OK;
}
bool SamplerSynth::readMaster(float* left, float* right) const {
    return d->ctx && zl_read_master(d->ctx, 0, left, right, framesInLastRender()) == ZL_OK;
}

void SamplerSynthVoice::refresh() {
    zl_ctx* ctx = m_synth->context();
    if (!ctx) return;
    std::vector<zl_voice_info> all(12 * 8);
    if (zl_read_voices(ctx, all.data(), (int32_t) all.size()) != ZL_OK) return;
    m_info = all[(size_t) m_channel * 8 + (size_t) m_slot];
    isPlaying = m_info.playing != 0;
}

// ---- ClipAudioSource ---------------------------------------------------------------------------------------------
class ClipAudioSource::Private {
public:
    int32_t clip = -1;
    std::string fileName, filePath;
    float duration = 0.0f;
    // lib/ClipAudioSource.cpp:63-82
    float startPositionInSeconds = 0, lengthInSeconds = -1, lengthInBeats = -1, volumeAbsolute = -1.0f;
    float pitchChange = 0, speedRatio = 1.0f, gainDb = 0, pan = 0.0f;
    double currentLeveldB = -400.0, prevLeveldB = -400.0;
    int id = 0;
    bool looping = true;
    ClipAudioSourcePositionsModel* positionsModel = nullptr;
    int slices = 0;
    std::vector<double> slicePositions;
    int sliceBaseMidiNote = 60, keyZoneStart = 0, keyZoneEnd = 127, rootNote = 60;
    ADSRParameters adsr;
    double firstPositionProgress = 0;
    void (*progressChangedCallback)(float) = nullptr;
    void (*audioLevelChangedCallback)(float) = nullptr;

    void push() {                                   // the device's clip record follows every change
        zl_ctx* ctx = SamplerSynth::instance()->context();
        if (!ctx || clip < 0) return;
        zl_clip_params p;
        memset(&p, 0, sizeof p);
        p.start_s = startPositionInSeconds; p.length_s = lengthInSeconds; p.length_beats = lengthInBeats; p.pan = pan;
        p.volume_abs = volumeAbsolute; p.attack = adsr.attack; p.decay = adsr.decay; p.sustain = adsr.sustain; p.release = adsr.release;
        p.root_note = rootNote; p.n_slices = (int32_t) slicePositions.size(); p.slice_pos = slicePositions.empty() ? nullptr : slicePositions.data();
        if (slicePositions.empty()) p.n_slices = 0;
        zl_clip_set_params(ctx, clip, &p);
    }
    void finish(ClipAudioSource* q, bool muted) {   // the tail of the reference constructor  lib/ClipAudioSource.cpp:150-205
        zl_ctx* ctx = SamplerSynth::instance()->context();
        zl_clip_params p;
        if (ctx && clip >= 0 && zl_clip_params_default(ctx, clip, &p) == ZL_OK) { duration = p.length_s; volumeAbsolute = p.volume_abs; }
        lengthInSeconds = duration;                                             // :158
        adsr = ADSRParameters(); adsr.attack = 0.0f; adsr.release = 0.05f;      // :164-168
        positionsModel = new ClipAudioSourcePositionsModel(q);
        if (muted) q->setVolume(-100.0f);                                       // :182-185
        SamplerSynth::instance()->registerClip(q);                              // :199
        q->setSlices(16);                                                       // :207
        push();
    }
};

ClipAudioSource::ClipAudioSource(tracktion_engine::Engine*, SyncTimer*, const char* filepath, bool muted) : d(new Private) {
    d->filePath = filepath ? filepath : "";
    const size_t slash = d->filePath.find_last_of('/');
    d->fileName = slash == std::string::npos ? d->filePath : d->filePath.substr(slash + 1);
    if (zl_ctx* ctx = SamplerSynth::instance()->context()) zl_clip_create_from_wav_file(ctx, d->filePath.c_str(), &d->clip);
    d->finish(this, muted);
}
ClipAudioSource::ClipAudioSource(const float* const* planar, int channels, int64_t frames, double sampleRate, bool muted) : d(new Private) {
    d->filePath = "memory"; d->fileName = "memory";
    if (zl_ctx* ctx = SamplerSynth::instance()->context()) zl_clip_create_from_pcm(ctx, planar, channels, frames, sampleRate, &d->clip);
    d->finish(this, muted);
}
ClipAudioSource::~ClipAudioSource() {               // lib/ClipAudioSource.cpp:207-219
    stop();
    SamplerSynth::instance()->unregisterClip(this);
    delete d->positionsModel;
    delete d;
}
int32_t ClipAudioSource::deviceClip() const { return d->clip; }

void ClipAudioSource::setProgressCallback(void (*functionPtr)(float)) { d->progressChangedCallback = functionPtr; }
void ClipAudioSource::syncProgress() {              // :225-240 (without the 100 ms rate limiter: the host decides when to call)
    double newPosition = d->startPositionInSeconds / getDuration();
    if (d->progressChangedCallback != nullptr && d->positionsModel && d->positionsModel->firstProgress() > -1.0f)
        newPosition = d->positionsModel->firstProgress();
    if (std::abs(d->firstPositionProgress - newPosition) > 0.001) {
        d->firstPositionProgress = newPosition;
        if (d->progressChangedCallback) d->progressChangedCallback((float) (d->firstPositionProgress * getDuration()));
    }
}
void ClipAudioSource::syncAudioLevel() {            // :88-113 (tracktion's level client reads silence here)
    d->prevLeveldB = d->currentLeveldB;
    d->currentLeveldB = std::max(gainToDecibels(d->positionsModel->peakGain()), -100.0);
    const auto prevLevel{decibelsToGain(d->prevLeveldB)};
    if (d->prevLeveldB > d->currentLeveldB) d->currentLeveldB = gainToDecibels(prevLevel * 0.94);
    if (std::abs(d->currentLeveldB - d->prevLeveldB) > 0.1 && d->audioLevelChangedCallback != nullptr)
        d->audioLevelChangedCallback((float) d->currentLeveldB);
}
void ClipAudioSource::setStartPosition(float s) { d->startPositionInSeconds = std::max(0.0f, s); d->push(); }   // :255-259
float ClipAudioSource::getStartPosition(int slice) const {                       // :261-268
    if (slice > -1 && slice < (int) d->slicePositions.size()) return (float) (d->startPositionInSeconds + (d->lengthInSeconds * d->slicePositions[(size_t) slice]));
    return d->startPositionInSeconds;
}
float ClipAudioSource::getStopPosition(int slice) const {                        // :270-277
    if (slice > -1 && slice + 1 < (int) d->slicePositions.size()) return (float) (d->startPositionInSeconds + (d->lengthInSeconds * d->slicePositions[(size_t) (slice + 1)]));
    return d->startPositionInSeconds + d->lengthInSeconds;
}
void ClipAudioSource::setLooping(bool looping) { d->looping = looping; }
bool ClipAudioSource::getLooping() const { return d->looping; }
void ClipAudioSource::setLength(float beat, int bpm) {                           // :352-360, lib/SyncTimer.cpp:180-183,936-939
    const uint64_t b = (uint64_t) std::min(200, std::max(50, bpm));              // BPM_MINIMUM / BPM_MAXIMUM
    const uint64_t ns = ((uint64_t) (beat * 96) * 60000000000ull) / (b * 96);
    d->lengthInSeconds = ns / (float) 1000000000;
    d->lengthInBeats = beat;
    d->push();
}
float ClipAudioSource::getLengthInBeats() const { return d->lengthInBeats; }
void ClipAudioSource::setPitch(float pitchChange, bool) { d->pitchChange = pitchChange; }
void ClipAudioSource::setSpeedRatio(float speedRatio, bool) { d->speedRatio = speedRatio; }
void ClipAudioSource::setGain(float db) { d->gainDb = db; }
void ClipAudioSource::setVolume(float vol) {                                     // :318-331
    d->volumeAbsolute = vol <= -40.0f ? 0.0f : decibelsToVolumeFaderPosition(vol);
    d->push();
}
void ClipAudioSource::setVolumeAbsolute(float vol) { d->volumeAbsolute = std::max(0.0f, std::min(vol, 1.0f)); d->push(); }   // :333-340
float ClipAudioSource::volumeAbsolute() const { return d->volumeAbsolute; }
void ClipAudioSource::setAudioLevelChangedCallback(void (*functionPtr)(float)) { d->audioLevelChangedCallback = functionPtr; }
void ClipAudioSource::play(bool loop, int midiChannel) {                         // :415-429
    ClipCommand* command = ClipCommand::channelCommand(this, midiChannel);
    command->midiNote = 60;
    command->changeVolume = true;
    command->volume = 1.0f;
    command->looping = loop;
    if (loop) command->stopPlayback = true;
    command->startPlayback = true;
    SamplerSynth::instance()->handleClipCommand(command, 0);
}
void ClipAudioSource::stop(int midiChannel) {                                    // :431-456
    SamplerSynth* synth = SamplerSynth::instance();
    if (midiChannel > -3) {
        ClipCommand* command = ClipCommand::channelCommand(this, midiChannel);
        command->midiNote = 60;
        command->stopPlayback = true;
        synth->handleClipCommand(command, 0);
    } else {
        ClipCommand* command = ClipCommand::noEffectCommand(this);
        command->stopPlayback = true;
        synth->handleClipCommand(command, 0);
        command = ClipCommand::effectedCommand(this);
        command->stopPlayback = true;
        synth->handleClipCommand(command, 0);
        for (int i = 0; i < 10; ++i) {
            command = ClipCommand::channelCommand(this, i);
            command->midiNote = 60;
            command->stopPlayback = true;
            synth->handleClipCommand(command, 0);
        }
    }
}
float ClipAudioSource::getDuration() { return d->duration; }
const char* ClipAudioSource::getFileName() const { return d->fileName.c_str(); }
const char* ClipAudioSource::getFilePath() const { return d->filePath.c_str(); }
int ClipAudioSource::id() const { return d->id; }
void ClipAudioSource::setId(int id) { d->id = id; }
float ClipAudioSource::audioLevel() const { return (float) d->currentLeveldB; }
double ClipAudioSource::position() const { return d->firstPositionProgress; }
ClipAudioSourcePositionsModel* ClipAudioSource::playbackPositionsModel() { return d->positionsModel; }
int ClipAudioSource::slices() const { return d->slices; }
void ClipAudioSource::setSlices(int slices) {                                    // :495-528
    if (d->slices != slices) {
        if (slices == 0) d->slicePositions.clear();
        else if (d->slices > slices) { while ((int) d->slicePositions.size() > slices) d->slicePositions.pop_back(); }
        else if (d->slices < slices) {
            double lastSlicePosition{0.0f};
            if (!d->slicePositions.empty()) lastSlicePosition = d->slicePositions.back();
            double positionIncrement{(1.0f - lastSlicePosition) / (slices - d->slices)};
            double newPosition{lastSlicePosition + positionIncrement};
            if (d->slicePositions.empty()) d->slicePositions.push_back(0.0f);
            while ((int) d->slicePositions.size() < slices) { d->slicePositions.push_back(newPosition); newPosition += positionIncrement; }
        }
        d->slices = slices;
        d->push();
    }
}
std::vector<double> ClipAudioSource::slicePositions() const { return d->slicePositions; }
void ClipAudioSource::setSlicePositions(const std::vector<double>& p) { if (d->slicePositions != p) { d->slicePositions = p; d->slices = (int) p.size(); d->push(); } }
double ClipAudioSource::slicePosition(int slice) const { return (slice > -1 && slice < (int) d->slicePositions.size()) ? d->slicePositions[(size_t) slice] : 0.0; }
void ClipAudioSource::setSlicePosition(int slice, float position) { if (slice > -1 && slice < (int) d->slicePositions.size()) { d->slicePositions[(size_t) slice] = position; d->push(); } }
int ClipAudioSource::sliceBaseMidiNote() const { return d->sliceBaseMidiNote; }
void ClipAudioSource::setSliceBaseMidiNote(int n) { d->sliceBaseMidiNote = n; }
int ClipAudioSource::sliceForMidiNote(int midiNote) const { return ((d->slices - (d->sliceBaseMidiNote % d->slices)) + midiNote) % d->slices; }   // :591-594
int ClipAudioSource::keyZoneStart() const { return d->keyZoneStart; }
void ClipAudioSource::setKeyZoneStart(int v) { d->keyZoneStart = v; }
int ClipAudioSource::keyZoneEnd() const { return d->keyZoneEnd; }
void ClipAudioSource::setKeyZoneEnd(int v) { d->keyZoneEnd = v; }
int ClipAudioSource::rootNote() const { return d->rootNote; }
void ClipAudioSource::setRootNote(int rootNote) { if (d->rootNote != rootNote) { d->rootNote = rootNote; d->push(); } }
float ClipAudioSource::pan() { return d->pan; }
void ClipAudioSource::setPan(float pan) { if (d->pan != pan) { d->pan = pan; d->push(); } }
float ClipAudioSource::adsrAttack() const { return d->adsr.attack; }
// :636-685: each setter builds a DEFAULT juce::ADSR::Parameters and sets one field - the other three fall back to 0.1/0.1/1.0/0.1
void ClipAudioSource::setADSRAttack(const float& v) { if (d->adsr.attack != v) { ADSRParameters p; p.attack = v; d->adsr = p; d->push(); } }
float ClipAudioSource::adsrDecay() const { return d->adsr.decay; }
void ClipAudioSource::setADSRDecay(const float& v) { if (d->adsr.decay != v) { ADSRParameters p; p.decay = v; d->adsr = p; d->push(); } }
float ClipAudioSource::adsrSustain() const { return d->adsr.sustain; }
void ClipAudioSource::setADSRSustain(const float& v) { if (d->adsr.sustain != v) { ADSRParameters p; p.sustain = v; d->adsr = p; d->push(); } }
float ClipAudioSource::adsrRelease() const { return d->adsr.release; }
void ClipAudioSource::setADSRRelease(const float& v) { if (d->adsr.release != v) { ADSRParameters p; p.release = v; d->adsr = p; d->push(); } }
const ADSRParameters& ClipAudioSource::adsrParameters() const { return d->adsr; }
void ClipAudioSource::setADSRParameters(const ADSRParameters& p) { d->adsr = p; d->push(); }

// ---- ClipAudioSourcePositionsModel ---------------------------------------------------------------------------------
float ClipAudioSourcePositionsModel::peakGain() const {
    zl_ctx* ctx = SamplerSynth::instance()->context();
    float g = 0.0f;
    if (ctx) zl_clip_peak_gain(ctx, m_clip->deviceClip(), &g, nullptr);
    return g;
}
double ClipAudioSourcePositionsModel::firstProgress() const {
    std::vector<float> g, p;
    positions(g, p);
    return p.empty() ? -1.0 : (double) p[0];
}
int ClipAudioSourcePositionsModel::rowCount() const { std::vector<float> g, p; positions(g, p); return (int) g.size(); }
void ClipAudioSourcePositionsModel::positions(std::vector<float>& gains, std::vector<float>& progress) const {
    gains.assign(32, 0.0f); progress.assign(32, 0.0f);
    int32_t n = 0;
    zl_ctx* ctx = SamplerSynth::instance()->context();
    if (ctx) zl_read_positions(ctx, m_clip->deviceClip(), gains.data(), progress.data(), 32, &n);
    gains.resize((size_t) n); progress.resize((size_t) n);
}

// ---- JackPassthrough  lib/JackPassthrough.cpp:160-230 ----------------------------------------------------------------
JackPassthrough::JackPassthrough(int channel) : m_bus(channel + 2) {}
zl_passthrough_params JackPassthrough::get() const {
    zl_passthrough_params p{1.0f, 1.0f, 1.0f, 0.0f, 0};
    if (zl_ctx* ctx = SamplerSynth::instance()->context()) zl_passthrough_get(ctx, 0, m_bus, &p);
    return p;
}
void JackPassthrough::set(const zl_passthrough_params& p) { if (zl_ctx* ctx = SamplerSynth::instance()->context()) zl_passthrough_set(ctx, 0, m_bus, &p); }
float JackPassthrough::dryAmount() const { return get().dry_amount; }
void JackPassthrough::setDryAmount(const float& v) { auto p = get(); if (p.dry_amount != v) { p.dry_amount = v; set(p); } }
float JackPassthrough::wetFx1Amount() const { return get().wet_fx1_amount; }
void JackPassthrough::setWetFx1Amount(const float& v) { auto p = get(); if (p.wet_fx1_amount != v) { p.wet_fx1_amount = v; set(p); } }
float JackPassthrough::wetFx2Amount() const { return get().wet_fx2_amount; }
void JackPassthrough::setWetFx2Amount(const float& v) { auto p = get(); if (p.wet_fx2_amount != v) { p.wet_fx2_amount = v; set(p); } }
float JackPassthrough::panAmount() const { return get().pan_amount; }
void JackPassthrough::setPanAmount(const float& v) { auto p = get(); if (p.pan_amount != v) { p.pan_amount = v; set(p); } }
bool JackPassthrough::muted() const { return get().muted != 0; }
void JackPassthrough::setMuted(const bool& v) { auto p = get(); if ((p.muted != 0) != v) { p.muted = v ? 1 : 0; set(p); } }

// ---- AudioLevels ---------------------------------------------------------------------------------------------------
namespace {
float convertTodbFS(float raw) {                    // lib/AudioLevels.cpp:330-341
    if (raw <= 0) return -200;
    const float fValue = 20 * log10f(raw);
    return fValue < -200 ? -200 : fValue;
}
bool writeFile(const std::string& path, const std::vector<unsigned char>& bytes) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) return false;
    const bool ok = fwrite(bytes.data(), 1, bytes.size(), f) == bytes.size();
    fclose(f);
    return ok;
}
bool endsWith(const std::string& s, const char* tail) { const size_t n = strlen(tail); return s.size() >= n && s.compare(s.size() - n, n, tail) == 0; }
}  // namespace

AudioLevels* AudioLevels::instance() { static AudioLevels* a = new AudioLevels(); return a; }
float AudioLevels::add(float db1, float db2) { return 10 * log10f(pow(10, db1 / 10) + pow(10, db2 / 10)); }   // lib/AudioLevels.cpp:234-236
void AudioLevels::timerCallback() {                 // lib/AudioLevels.cpp:347-412
    SamplerSynth* synth = SamplerSynth::instance();
    zl_ctx* ctx = synth->context();
    if (!ctx) return;
    // the reference scans the most recent period's buffer once (bufferReadSize is cleared after the scan): a tick without
    // a new render only decays
    const int block = synth->renderSerial() != m_scannedSerial ? synth->blocksInLastRender() - 1 : -1;
    m_scannedSerial = synth->renderSerial();
    zl_levels lv;
    if (zl_levels_tick(ctx, 0, 0, block, 1, &lv) == ZL_OK) {       // "system playback": the global un-effected channel's dry port
        playbackA = lv.db_a; playbackB = lv.db_b; playback = lv.db_sum;
        playbackAHold = convertTodbFS(lv.hold_a); playbackBHold = convertTodbFS(lv.hold_b);
    }
    for (int t = 0; t < 10; ++t)
        if (zl_levels_tick(ctx, 0, t + 2, block, 0, &lv) == ZL_OK) { channelsA[t] = lv.db_a; channelsB[t] = lv.db_b; m_levels[(size_t) t] = lv.db_sum; }
}
std::vector<float> AudioLevels::channels() const { return m_levels; }
void AudioLevels::setChannelToRecord(int channel, bool shouldRecord) { if (channel > -1 && channel < 10) m_recordChannel[channel] = shouldRecord; }
void AudioLevels::setChannelFilenamePrefix(int channel, const std::string& p) { if (channel > -1 && channel < 10) m_prefix[channel] = p; }
void AudioLevels::setRecordGlobalPlayback(bool shouldRecord) { m_recordGlobal = shouldRecord; }
void AudioLevels::setGlobalPlaybackFilenamePrefix(const std::string& p) { m_globalPrefix = p; }
void AudioLevels::startRecording() {                // lib/AudioLevels.cpp:514-560, DiskWriter :35-119: 16-bit stereo WAV
    zl_ctx* ctx = SamplerSynth::instance()->context();
    if (!ctx) return;
    auto bounce = [&](int bus, const std::string& path) {
        uint64_t size = 0;
        zl_bounce_wav(ctx, 0, bus, ZL_BUS_DRY, 16, nullptr, 0, &size);
        if (!size) return;
        std::vector<unsigned char> img((size_t) size);
        if (zl_bounce_wav(ctx, 0, bus, ZL_BUS_DRY, 16, img.data(), size, &size) == ZL_OK) writeFile(path, img);
    };
    const std::string timestamp = "recording";      // the reference appends an ISO date; the host names the take here
    if (m_recordGlobal) bounce(-1, endsWith(m_globalPrefix, ".wav") ? m_globalPrefix : m_globalPrefix + "-" + timestamp + ".wav");
    for (int t = 0; t < 10; ++t)
        if (m_recordChannel[t]) bounce(t + 2, m_prefix[t] + "-" + timestamp + ".wav");
}

// ---- WaveFormItem  lib/WaveFormItem.cpp:96-143 ---------------------------------------------------------------------
void WaveFormItem::setSource(const std::string& source) {
    if (m_source == source) return;
    m_source = source;
    m_owned.reset(new ClipAudioSource(nullptr, nullptr, source.c_str(), true));
    m_clip = m_owned.get();
    m_start = 0; m_end = length();
}
void WaveFormItem::setSource(ClipAudioSource* clip) { m_owned.reset(); m_clip = clip; m_source = clip ? clip->getFilePath() : ""; m_start = 0; m_end = length(); }
double WaveFormItem::length() const { return m_clip ? (double) m_clip->getDuration() : 0.0; }
bool WaveFormItem::paint(int width, int height, std::vector<uint8_t>& mask) const {
    zl_ctx* ctx = SamplerSynth::instance()->context();
    if (!ctx || !m_clip || m_clip->deviceClip() < 0 || width < 1 || height < 1) return false;
    mask.assign((size_t) width * (size_t) height, 0);
    return zl_thumb_raster(ctx, m_clip->deviceClip(), 0, m_start, std::min(m_end, length()), width, height, 1.0f, mask.data()) == ZL_OK;
}
bool WaveFormItem::paintARGB(int width, int height, std::vector<uint32_t>& pixels) const {
    std::vector<uint8_t> mask;
    if (!paint(width, height, mask)) return false;
    pixels.resize(mask.size());
    for (size_t i = 0; i < mask.size(); ++i) pixels[i] = mask[i] ? m_color : 0u;
    return true;
}

// ---- SyncTimer (row n2) --------------------------------------------------------------------------------------------
class SyncTimer::Private {
public:
    zl_sched* sched = nullptr;
    double rate = 48000.0;
    uint64_t bpm = 120;
    bool running = false;
    mutable uint64_t playhead = 0, playheadUsecs = 0, subbeatUsecs = 0;
    uint64_t sent = 0, periods = 0;
    std::map<int32_t, ClipAudioSource*> clips;                       // device clip id -> the object the host scheduled
    std::vector<zl_block_clock> clocks;
    std::vector<zl_sched_event> events;
    void refresh() const { zl_sched_state(sched, &playhead, &playheadUsecs, &subbeatUsecs, nullptr, nullptr, nullptr); }
    static void toC(const ClipCommand& c, zl_clip_command& k) {
        zl_clip_command_default(&k);
        k.clip = c.clip ? c.clip->deviceClip() : -1; k.midi_note = c.midiNote; k.midi_channel = c.midiChannel;
        k.start_playback = c.startPlayback; k.stop_playback = c.stopPlayback; k.change_slice = c.changeSlice; k.slice = c.slice;
        k.change_looping = c.changeLooping; k.looping = c.looping; k.change_pitch = c.changePitch; k.pitch_change = c.pitchChange;
        k.change_speed = c.changeSpeed; k.speed_ratio = c.speedRatio; k.change_gain_db = c.changeGainDb; k.gain_db = c.gainDb;
        k.change_volume = c.changeVolume; k.volume = c.volume;
    }
    void fromC(const zl_clip_command& k, ClipCommand& c) const {
        auto it = clips.find(k.clip);
        c.clip = it == clips.end() ? nullptr : it->second; c.midiNote = k.midi_note; c.midiChannel = k.midi_channel;
        c.startPlayback = k.start_playback != 0; c.stopPlayback = k.stop_playback != 0; c.changeSlice = k.change_slice != 0; c.slice = k.slice;
        c.changeLooping = k.change_looping != 0; c.looping = k.looping != 0; c.changePitch = k.change_pitch != 0; c.pitchChange = k.pitch_change;
        c.changeSpeed = k.change_speed != 0; c.speedRatio = k.speed_ratio; c.changeGainDb = k.change_gain_db != 0; c.gainDb = k.gain_db;
        c.changeVolume = k.change_volume != 0; c.volume = k.volume;
    }
};

SyncTimer* SyncTimer::instance() {
    static SyncTimer* instance{nullptr};
    if (!instance) instance = new SyncTimer();
    return instance;
}
SyncTimer::SyncTimer(double sampleRate, int bufferSize) : d(new Private) {
    d->rate = sampleRate;
    zl_sched_create(sampleRate, bufferSize, &d->sched);
    d->refresh();
}
SyncTimer::~SyncTimer() { if (d->sched) zl_sched_destroy(d->sched); delete d; }
zl_sched* SyncTimer::scheduler() const { return d->sched; }

void SyncTimer::queueClipToStart(ClipAudioSource* clip) { queueClipToStartOnChannel(clip, -1); }       // lib/SyncTimer.cpp:862-868
void SyncTimer::queueClipToStop(ClipAudioSource* clip) { queueClipToStopOnChannel(clip, -1); }
void SyncTimer::queueClipToStartOnChannel(ClipAudioSource* clip, int midiChannel) {
    if (!clip) return;
    d->clips[clip->deviceClip()] = clip;
    zl_sched_queue_start(d->sched, clip->deviceClip(), midiChannel);
}
void SyncTimer::queueClipToStopOnChannel(ClipAudioSource* clip, int midiChannel) {
    if (!clip) return;
    d->clips[clip->deviceClip()] = clip;
    zl_sched_queue_stop(d->sched, clip->deviceClip(), midiChannel);
}
void SyncTimer::start(int bpm) { d->bpm = (uint64_t) bpm; zl_sched_start(d->sched, bpm); d->running = true; }
void SyncTimer::stop() { zl_sched_stop(d->sched); d->running = false; }
int SyncTimer::getInterval(int bpm) { return 60000 / (bpm * 96); }                                     // :928-931
static uint64_t boundBpm(uint64_t bpm) { return bpm < 50 ? 50 : (bpm > 200 ? 200 : bpm); }             // BPM_MINIMUM / BPM_MAXIMUM  :28-29
float SyncTimer::subbeatCountToSeconds(uint64_t bpm, uint64_t beats) const {                           // :936-939, SyncTimerThread::subbeatCountToNanoseconds
    return (float) ((beats * 60000000000ull) / (boundBpm(bpm) * 96ull)) / (float) 1000000000;
}
uint64_t SyncTimer::secondsToSubbeatCount(uint64_t bpm, float seconds) const {                         // :941-944: a float quotient, truncated
    const uint64_t nanoseconds = (uint64_t) floor(seconds * (float) 1000000000);
    const float count = nanoseconds / (60000000000ull / (boundBpm(bpm) * 96ull));
    return (uint64_t) count;
}
int SyncTimer::getMultiplier() { return 96; }
uint64_t SyncTimer::getBpm() const { return d->bpm; }
void SyncTimer::setBpm(uint64_t bpm) { if (bpm != d->bpm) { d->bpm = bpm; zl_sched_set_bpm(d->sched, (int32_t) bpm); } }
uint64_t SyncTimer::scheduleAheadAmount() const { uint64_t v = 0; zl_sched_state(d->sched, nullptr, nullptr, nullptr, nullptr, nullptr, &v); return v; }
int SyncTimer::beat() const { int32_t v = 0; zl_sched_state(d->sched, nullptr, nullptr, nullptr, nullptr, &v, nullptr); return v; }
uint64_t SyncTimer::cumulativeBeat() const { uint64_t v = 0; zl_sched_state(d->sched, nullptr, nullptr, nullptr, &v, nullptr, nullptr); return v; }
const uint64_t& SyncTimer::jackPlayhead() const { d->refresh(); return d->playhead; }
const uint64_t& SyncTimer::jackPlayheadUsecs() const { d->refresh(); return d->playheadUsecs; }
const uint64_t& SyncTimer::jackSubbeatLengthInMicroseconds() const { d->refresh(); return d->subbeatUsecs; }
bool SyncTimer::timerRunning() { return d->running; }
ClipCommand* SyncTimer::getClipCommand() { return new ClipCommand(); }                                 // the reference hands out pool entries (:1115-1128)
void SyncTimer::deleteClipCommand(ClipCommand* command) { delete command; }

void SyncTimer::scheduleClipCommand(ClipCommand* command, uint64_t delay) {
    if (!command) return;
    if (command->clip) d->clips[command->clip->deviceClip()] = command->clip;
    zl_clip_command k;
    Private::toC(*command, k);
    zl_sched_schedule(d->sched, &k, delay);
    delete command;                                               // merged into a waiting command or copied into the ring
}
void SyncTimer::scheduleTimerCommand(uint64_t delay, int operation, int parameter1, int parameter2, int) {
    if (operation == 8) zl_sched_schedule_channel_enabled(d->sched, delay, parameter1, parameter2 != 0);   // TimerCommand::SamplerChannelEnabledStateOperation
}

bool SyncTimer::process(uint32_t nframes, uint64_t currentUsecs, uint64_t nextUsecs) {
    d->clocks.assign(1, zl_block_clock{});
    d->clocks[0].current_usecs = currentUsecs; d->clocks[0].next_usecs = nextUsecs;
    SamplerSynth* synth = SamplerSynth::instance();
    if (!synth->isInitialized()) return false;
    d->events.resize(4096);
    int32_t n = 0;
    int rc = zl_sched_advance(d->sched, 1, (int32_t) nframes, d->clocks.data(), d->events.data(), (int32_t) d->events.size(), &n);
    if (rc != ZL_OK) return false;
    for (int32_t i = 0; i < n; ++i) {
        const zl_sched_event& ev = d->events[(size_t) i];
        if (ev.kind == ZL_SCHED_CLIP_COMMAND) {
            zl_command(synth->context(), 0, 0, ev.tick, &ev.cmd);
            ++d->sent;
            if (clipCommandSent) { ClipCommand c; d->fromC(ev.cmd, c); clipCommandSent(&c); }
        } else synth->setChannelEnabled(ev.channel, ev.enabled != 0);
    }
    ++d->periods;
    return synth->process(nframes, d->clocks[0]);
}

bool SyncTimer::render(int nblocks, uint32_t nframes, uint64_t firstUsecs) {
    SamplerSynth* synth = SamplerSynth::instance();
    if (!synth->isInitialized() || nblocks < 1) return false;
    const uint64_t period = (uint64_t) llround((double) nframes * 1.0e6 / d->rate);
    d->clocks.assign((size_t) nblocks, zl_block_clock{});
    for (int b = 0; b < nblocks; ++b) {
        d->clocks[(size_t) b].current_usecs = firstUsecs + (uint64_t) b * period;
        d->clocks[(size_t) b].next_usecs = firstUsecs + (uint64_t) (b + 1) * period;
    }
    d->events.resize(65536);
    int32_t n = 0;
    if (zl_sched_advance(d->sched, nblocks, (int32_t) nframes, d->clocks.data(), d->events.data(), (int32_t) d->events.size(), &n) != ZL_OK) return false;
    d->periods += (uint64_t) nblocks;
    // a channel switched on or off takes effect from its own period on: the batch is rendered in segments that start there;
    // SamplerSynth keeps only the last segment's ports, so hosts that switch channels mid-batch read back per segment
    // (process()) - the bounce of a whole batch with switches is what tests/libref.py does through the C ABI
    int lo = 0;
    bool ok = true;
    int32_t i = 0;
    while (lo < nblocks && ok) {
        int hi = nblocks;
        for (int32_t j = i; j < n; ++j)
            if (d->events[(size_t) j].kind != ZL_SCHED_CLIP_COMMAND && d->events[(size_t) j].block > lo) { hi = d->events[(size_t) j].block; break; }
        for (; i < n && d->events[(size_t) i].block < hi; ++i) {
            const zl_sched_event& ev = d->events[(size_t) i];
            if (ev.kind == ZL_SCHED_CLIP_COMMAND) {
                zl_command(synth->context(), 0, ev.block - lo, ev.tick, &ev.cmd);
                ++d->sent;
                if (clipCommandSent) { ClipCommand c; d->fromC(ev.cmd, c); clipCommandSent(&c); }
            } else synth->setChannelEnabled(ev.channel, ev.enabled != 0);
        }
        ok = synth->render(hi - lo, nframes, d->clocks.data() + lo);
        lo = hi;
    }
    return ok;
}
uint64_t SyncTimer::commandsSent() const { return d->sent; }
uint64_t SyncTimer::periodsProcessed() const { return d->periods; }
