// oracle/ref/libzl_ref.cpp — TEST INFRASTRUCTURE.  Drives the reference's OWN hot-path sources, compiled in place from
// /root/reference/lib by oracle/Makefile:
//
//   lib/SamplerSynthVoice.cpp             SamplerSynthVoice::process / startNote / stopNote / setCurrentCommand   (rows a1, a3)
//   lib/SamplerSynth.cpp                  SamplerChannel::process / handleCommand, the command ring, voice allocation (a5, a6)
//   lib/JackPassthrough.cpp               JackPassthroughPrivate::process                                        (a9)
//   lib/ClipAudioSourcePositionsModel.cpp the positions model, peakGain with its hysteresis                      (a14)
//   juce_ADSR.h, juce_Synthesiser.cpp     through the reference's own JUCE objects already in oracle/_ref        (a4)
//
// What those files need around them is NOT the reference's arithmetic and is supplied here or in shim/:
//   * Qt (containers, strings, the object tree; signals get empty bodies below where moc would have generated them),
//   * JACK: an in-process stand-in - ports are float buffers, the test calls the registered process callbacks in client
//     order with the cycle times it chooses (jack_get_cycle_times),
//   * SyncTimer: the reference's own lib/SyncTimer.cpp (see synctimer_ref.cpp); the voice scenarios inject the playhead values
//     the voice reads, the scheduler scenarios run its step ring for real (row n2),
//   * ClipAudioSource: lib/ClipAudioSource.cpp is a tracktion_engine Edit wrapper; the voice only calls its getters, so the
//     class declared by lib/ClipAudioSource.h gets a parameter-holder implementation here (getStartPosition /
//     getStopPosition restated from lib/ClipAudioSource.cpp:261-277 - row a8 stays a restatement),
//   * SamplerSynthSound: lib/SamplerSynthSound.cpp reads the file through tracktion's AudioFile; here the buffer is filled
//     from the caller's PCM the way loadSoundData does (AudioBuffer<float>(min(2, channels), length), lib/SamplerSynthSound.cpp
//     :40-46) and the getters are the file's own one-liners (:76-112).
//
// Nothing under libzl_b200/ links, loads or calls this file.
#include "JUCEHeaders.h"
#include "ClipAudioSource.h"
#include "ClipAudioSourcePositionsModel.h"
#include "ClipCommand.h"
#include "JackPassthrough.h"
#include "SamplerSynth.h"
#include "SamplerSynthSound.h"
#include "SamplerSynthVoice.h"
#include "SyncTimer.h"
#include "libzl.h"

#include <jack/jack.h>
#include <jack/midiport.h>

#include <cstdarg>
#include <cstring>
#include <map>
#include <string>
#include <vector>

extern "C" long long zl_ref_shim_now_msecs = 1000000;

// ---------------------------------------------------------------------------------------------------------------------
// JACK stand-in
struct _jack_port {
    std::string name; unsigned long flags = 0; bool midi = false;
    std::vector<float> buf;              // nframes + 1: quirk Q2 writes index nframes (lib/SamplerSynthVoice.cpp:218-221)
    _jack_port* source = nullptr;        // an input port with one connection reads its source's buffer
    std::vector<_jack_port*> mix;        // more than one: the sum of the sources in connection order (what jackd's mixdown does)
};
struct _jack_client {
    std::string name; JackProcessCallback cb = nullptr; void* arg = nullptr; bool active = false;
    std::vector<_jack_port*> ports;
};
namespace {
struct FakeJack {
    std::vector<_jack_client*> clients;
    jack_nframes_t sampleRate = 48000, bufferSize = 1024, currentFrames = 0;
    jack_time_t currentUsecs = 0, nextUsecs = 0;
    float periodUsecs = 0;
};
FakeJack gJack;
_jack_port* findPort(const char* full) {
    for (auto* c : gJack.clients)
        for (auto* p : c->ports)
            if (c->name + ":" + p->name == full) return p;
    return nullptr;
}
}  // namespace
extern "C" {
jack_client_t* jack_client_open(const char* name, jack_options_t, jack_status_t* status, ...) {
    auto* c = new _jack_client; c->name = name; gJack.clients.push_back(c);
    if (status) *status = (jack_status_t) 0;
    return c;
}
int jack_client_close(jack_client_t* c) {
    for (auto it = gJack.clients.begin(); it != gJack.clients.end(); ++it) if (*it == c) { gJack.clients.erase(it); break; }
    for (auto* p : c->ports) delete p;
    delete c; return 0;
}
int jack_set_process_callback(jack_client_t* c, JackProcessCallback cb, void* arg) { c->cb = cb; c->arg = arg; return 0; }
int jack_set_xrun_callback(jack_client_t*, JackXRunCallback, void*) { return 0; }
int jack_set_buffer_size_callback(jack_client_t*, JackBufferSizeCallback, void*) { return 0; }
int jack_set_sample_rate_callback(jack_client_t*, JackSampleRateCallback, void*) { return 0; }
int jack_set_port_registration_callback(jack_client_t*, JackPortRegistrationCallback, void*) { return 0; }
void jack_on_shutdown(jack_client_t*, JackShutdownCallback, void*) {}
jack_port_t* jack_port_register(jack_client_t* c, const char* name, const char* type, unsigned long flags, unsigned long) {
    auto* p = new _jack_port; p->name = name; p->flags = flags; p->midi = strcmp(type, JACK_DEFAULT_MIDI_TYPE) == 0;
    p->buf.assign((size_t) gJack.bufferSize + 8, 0.0f);
    c->ports.push_back(p); return p;
}
int jack_activate(jack_client_t* c) { c->active = true; return 0; }
int jack_deactivate(jack_client_t* c) { c->active = false; return 0; }
int jack_connect(jack_client_t*, const char* src, const char* dst) {
    _jack_port* s = findPort(src); _jack_port* d = findPort(dst);
    if (!s || !d) return -1;                       // system:playback_* and the MIDI router do not exist here: the reference only logs it
    d->source = s; return 0;
}
int jack_disconnect(jack_client_t*, const char*, const char* dst) { if (_jack_port* d = findPort(dst)) d->source = nullptr; return 0; }
void* jack_port_get_buffer(jack_port_t* p, jack_nframes_t nframes) {
    if (!p->mix.empty()) {
        for (jack_nframes_t f = 0; f < nframes; ++f) {
            float acc = p->mix[0]->buf[f];
            for (size_t k = 1; k < p->mix.size(); ++k) acc += p->mix[k]->buf[f];
            p->buf[f] = acc;
        }
        return p->buf.data();
    }
    return (p->source ? p->source : p)->buf.data();
}
const char* jack_port_name(const jack_port_t* p) { return p->name.c_str(); }
int jack_port_flags(const jack_port_t* p) { return (int) p->flags; }
jack_port_t* jack_port_by_id(jack_client_t*, jack_port_id_t) { return nullptr; }
jack_port_t* jack_port_by_name(jack_client_t*, const char* name) { return findPort(name); }
const char** jack_get_ports(jack_client_t*, const char*, const char*, unsigned long) { return nullptr; }
void jack_free(void*) {}
int jack_get_cycle_times(const jack_client_t*, jack_nframes_t* cf, jack_time_t* cu, jack_time_t* nu, float* pu) {
    *cf = gJack.currentFrames; *cu = gJack.currentUsecs; *nu = gJack.nextUsecs; *pu = gJack.periodUsecs; return 0;
}
float jack_cpu_load(jack_client_t*) { return 0.0f; }
jack_nframes_t jack_get_sample_rate(jack_client_t*) { return gJack.sampleRate; }
jack_nframes_t jack_get_buffer_size(jack_client_t*) { return gJack.bufferSize; }
jack_time_t jack_get_time(void) { return gJack.currentUsecs; }
jack_nframes_t jack_frame_time(const jack_client_t*) { return gJack.currentFrames; }
jack_nframes_t jack_last_frame_time(const jack_client_t*) { return gJack.currentFrames; }
jack_time_t jack_frames_to_time(const jack_client_t*, jack_nframes_t) { return gJack.currentUsecs; }
jack_transport_state_t jack_transport_query(const jack_client_t*, jack_position_t*) { return JackTransportStopped; }
int jack_set_latency_callback(jack_client_t*, JackLatencyCallback, void*) { return 0; }
void jack_port_get_latency_range(jack_port_t*, jack_latency_callback_mode_t, jack_latency_range_t* range) { range->min = 0; range->max = 0; }
uint32_t jack_midi_get_event_count(void*) { return 0; }
int jack_midi_event_get(jack_midi_event_t*, void*, uint32_t) { return -1; }
void jack_midi_clear_buffer(void*) {}
int jack_midi_event_write(void*, jack_nframes_t, const jack_midi_data_t*, size_t) { return 0; }
jack_midi_data_t* jack_midi_event_reserve(void*, jack_nframes_t, size_t) { return nullptr; }
}

// ---------------------------------------------------------------------------------------------------------------------
// SyncTimer is the reference's own lib/SyncTimer.cpp (compiled through synctimer_ref.cpp, which also owns these hooks)
void zlref_timer_inject(unsigned long long playhead, unsigned long long playhead_usecs, unsigned long long subbeat_usecs);
void zlref_timer_handle(ClipCommand* c, unsigned long long tick);
void zlref_timer_enable(int channel, bool enabled);
void zlref_timer_boot();                            // initJuce()'s "syncTimer = SyncTimer::instance()" (lib/libzl.cpp:388-389)

// ---------------------------------------------------------------------------------------------------------------------
// ClipAudioSource: the class lib/ClipAudioSource.h declares, as the parameter holder the voice and the sound query
namespace {
struct ClipData {
    std::vector<std::vector<float>> pcm;       // the "file": planar channels
    double rate = 48000.0;
    float startPositionInSeconds = 0.0f, lengthInSeconds = -1.0f, lengthInBeats = -1.0f, volumeAbsolute = -1.0f, pan = 0.0f;
    int rootNote = 60, id = 0;
    juce::ADSR adsr;
    QList<double> slicePositionsCache;
    ClipAudioSourcePositionsModel* positionsModel = nullptr;
    std::string path = "memory";
};
std::map<const ClipAudioSource*, ClipData*> gClipData;     // the class keeps its d pointer private; the driver goes through this
}  // namespace
class ClipAudioSource::Private : public ClipData {};
ClipAudioSource::ClipAudioSource(tracktion_engine::Engine*, SyncTimer*, const char*, bool, QObject* parent) : QObject(parent), d(new Private) {
    d->positionsModel = new ClipAudioSourcePositionsModel(this);
    gClipData[this] = d;
}
ClipAudioSource::~ClipAudioSource() { gClipData.erase(this); delete d->positionsModel; delete d; }
float ClipAudioSource::getStartPosition(int slice) const {                    // lib/ClipAudioSource.cpp:261-268
    if (slice > -1 && slice < d->slicePositionsCache.length()) return d->startPositionInSeconds + (d->lengthInSeconds * d->slicePositionsCache[slice]);
    return d->startPositionInSeconds;
}
float ClipAudioSource::getStopPosition(int slice) const {                     // :270-277
    if (slice > -1 && slice + 1 < d->slicePositionsCache.length()) return d->startPositionInSeconds + (d->lengthInSeconds * d->slicePositionsCache[slice + 1]);
    return d->startPositionInSeconds + d->lengthInSeconds;
}
float ClipAudioSource::getLengthInBeats() const { return d->lengthInBeats; }
float ClipAudioSource::volumeAbsolute() const { return d->volumeAbsolute; }
float ClipAudioSource::getDuration() { return (float) ((double) d->pcm[0].size() / d->rate); }   // edit->getLength(): the file's length in seconds
const char* ClipAudioSource::getFilePath() const { return d->path.c_str(); }
ClipAudioSourcePositionsModel* ClipAudioSource::playbackPositionsModel() { return d->positionsModel; }
int ClipAudioSource::rootNote() const { return d->rootNote; }
float ClipAudioSource::pan() { return d->pan; }
const juce::ADSR::Parameters& ClipAudioSource::adsrParameters() const { return d->adsr.getParameters(); }

// SamplerSynthSound over that holder
class SamplerSynthSoundPrivate {
public:
    bool isValid = false;
    std::unique_ptr<juce::AudioBuffer<float>> data;
    int length = 0;
    double sourceSampleRate = 0.0;
    ClipAudioSource* clip = nullptr;
};
namespace { ClipData* gLastClipPrivate = nullptr; }
SamplerSynthSound::SamplerSynthSound(ClipAudioSource* clip) : juce::SynthesiserSound(), d(new SamplerSynthSoundPrivate) {
    d->clip = clip;
    ClipData* cp = gLastClipPrivate;                                           // loadSoundData  lib/SamplerSynthSound.cpp:28-59
    d->sourceSampleRate = cp->rate;
    if (d->sourceSampleRate > 0 && !cp->pcm.empty() && !cp->pcm[0].empty()) {
        d->length = (int) cp->pcm[0].size();
        d->data.reset(new juce::AudioBuffer<float>(juce::jmin(2, (int) cp->pcm.size()), d->length));
        for (int ch = 0; ch < d->data->getNumChannels(); ++ch) memcpy(d->data->getWritePointer(ch), cp->pcm[(size_t) ch].data(), sizeof(float) * (size_t) d->length);
        d->isValid = true;
    }
}
SamplerSynthSound::~SamplerSynthSound() { delete d; }
ClipAudioSource* SamplerSynthSound::clip() const { return d->clip; }
bool SamplerSynthSound::isValid() const { return d->isValid; }
juce::AudioBuffer<float>* SamplerSynthSound::audioData() const noexcept { return d->data.get(); }
int SamplerSynthSound::length() const { return d->length; }
int SamplerSynthSound::startPosition(int slice) const { return d->clip->getStartPosition(slice) * d->sourceSampleRate; }   // :96-99
int SamplerSynthSound::stopPosition(int slice) const { return d->clip->getStopPosition(slice) * d->sourceSampleRate; }     // :101-104
int SamplerSynthSound::rootMidiNote() const { return d->clip->rootNote(); }
double SamplerSynthSound::sourceSampleRate() const { return d->sourceSampleRate; }

// signals: where moc would have generated emitters
void ClipAudioSourcePositionsModel::peakGainChanged() {}
void ClipAudioSourcePositionsModel::positionIDCreated(void*, qint64) {}
void JackPassthrough::dryAmountChanged() {}
void JackPassthrough::wetFx1AmountChanged() {}
void JackPassthrough::wetFx2AmountChanged() {}
void JackPassthrough::panAmountChanged() {}
void JackPassthrough::mutedChanged() {}

// ---------------------------------------------------------------------------------------------------------------------
// C API for the tests (ctypes)
namespace {
struct World {
    std::vector<ClipAudioSource*> clips;
    std::vector<JackPassthrough*> passthrough;      // one per SamplerChannel, wired to its left_out / right_out
    bool ready = false;
    int nframesMax = 0;
};
World gWorld;
_jack_client* channelClient(int bus) {              // lib/SamplerSynth.cpp:258-270 names
    std::string n = bus == 0 ? "SamplerSynth-global-uneffected" : bus == 1 ? "SamplerSynth-global-effected" : "SamplerSynth-channel_" + std::to_string(bus - 1);
    for (auto* c : gJack.clients) if (c->name == n) return c;
    return nullptr;
}
_jack_client* passthroughClient(int bus) {
    for (auto* c : gJack.clients) if (c->name == "ZLRefPassthrough" + std::to_string(bus)) return c;
    return nullptr;
}
_jack_port* portOf(_jack_client* c, const char* name) { for (auto* p : c->ports) if (p->name == name) return p; return nullptr; }
#define ZLREF __attribute__((visibility("default")))
}  // namespace

extern "C" {

// SamplerSynth::initialize (12 SamplerChannels x 8 voices, lib/SamplerSynth.cpp:254-278) plus one JackPassthrough per channel.
// One world per process (the reference's objects are singletons); max_frames bounds the JACK period.
ZLREF int zlref_init(double sample_rate, int max_frames) {
    if (gWorld.ready) return (gJack.sampleRate == (jack_nframes_t) sample_rate && max_frames <= gWorld.nframesMax) ? 0 : -1;
    gJack.sampleRate = (jack_nframes_t) sample_rate; gJack.bufferSize = (jack_nframes_t) max_frames; gWorld.nframesMax = max_frames;
    zlref_timer_boot();                             // the order of initJuce(): the timer, then SamplerSynth::initialize
    SamplerSynth::instance()->initialize(nullptr);
    for (int b = 0; b < 12; ++b) {
        auto* pt = new JackPassthrough(QString(("ZLRefPassthrough" + std::to_string(b)).c_str()));
        gWorld.passthrough.push_back(pt);
        _jack_client* src = channelClient(b); _jack_client* dst = passthroughClient(b);
        if (!src || !dst) return -2;
        portOf(dst, "inputLeft")->source = portOf(src, "left_out");
        portOf(dst, "inputRight")->source = portOf(src, "right_out");
    }
    gWorld.ready = true;
    return 0;
}

// ClipAudioSource_new + SamplerSynth::registerClip over in-memory PCM; returns the clip index
ZLREF int zlref_clip_new(const float* const* planar, int nch, long long n, double rate) {
    auto* clip = new ClipAudioSource(nullptr, SyncTimer::instance(), "memory");
    ClipData* d = gClipData[clip];
    d->pcm.resize((size_t) nch);
    for (int ch = 0; ch < nch; ++ch) d->pcm[(size_t) ch].assign(planar[ch], planar[ch] + n);
    d->rate = rate;
    d->lengthInSeconds = (float) ((double) n / rate);                        // lib/ClipAudioSource.cpp:158
    d->adsr.setSampleRate(rate);
    d->adsr.setParameters(juce::ADSR::Parameters(0.0f, 0.1f, 1.0f, 0.05f));   // :164-168
    d->volumeAbsolute = std::exp((-3.0f - 6.0f) * (1.0f / 20.0f));
    d->slicePositionsCache.clear();
    for (int i = 0; i < 16; ++i) d->slicePositionsCache << (double) i * (1.0 / 16.0);
    d->id = (int) gWorld.clips.size();
    gLastClipPrivate = d;
    SamplerSynth::instance()->registerClip(clip);
    gWorld.clips.push_back(clip);
    return d->id;
}
ZLREF int zlref_clip_set(int clip, float start_s, float length_s, float length_beats, float pan, float volume_abs,
                         float attack, float decay, float sustain, float release, int root_note, const double* slices, int n_slices) {
    if (clip < 0 || clip >= (int) gWorld.clips.size()) return -1;
    ClipData* d = gClipData[gWorld.clips[(size_t) clip]];
    d->startPositionInSeconds = start_s; d->lengthInSeconds = length_s; d->lengthInBeats = length_beats; d->pan = pan;
    d->volumeAbsolute = volume_abs; d->rootNote = root_note;
    d->adsr.setParameters(juce::ADSR::Parameters(attack, decay, sustain, release));
    d->slicePositionsCache.clear();
    for (int i = 0; i < n_slices; ++i) d->slicePositionsCache << slices[i];
    return 0;
}

// SyncTimer -> SamplerSynth::handleClipCommand(command, tick): lands in the channel's ring, applied by its next process()
ZLREF int zlref_command(int clip, int midi_note, int midi_channel, int start_playback, int stop_playback, int change_slice, int slice,
                        int change_looping, int looping, int change_volume, float volume, unsigned long long tick) {
    if (clip < 0 || clip >= (int) gWorld.clips.size()) return -1;
    ClipCommand* c = SyncTimer::instance()->getClipCommand();
    c->clip = gWorld.clips[(size_t) clip]; c->midiNote = midi_note; c->midiChannel = midi_channel;
    c->startPlayback = start_playback; c->stopPlayback = stop_playback; c->changeSlice = change_slice; c->slice = slice;
    c->changeLooping = change_looping; c->looping = looping; c->changeVolume = change_volume; c->volume = volume;
    zlref_timer_handle(c, tick);
    return 0;
}
ZLREF int zlref_channel_enabled(int midi_channel, int enabled) { zlref_timer_enable(midi_channel, enabled != 0); return 0; }
ZLREF int zlref_passthrough_set(int bus, float dry, float fx1, float fx2, float pan, int muted) {
    if (bus < 0 || bus >= 12) return -1;
    JackPassthrough* p = gWorld.passthrough[(size_t) bus];
    p->setDryAmount(dry); p->setWetFx1Amount(fx1); p->setWetFx2Amount(fx2); p->setPanAmount(pan); p->setMuted(muted != 0);
    return 0;
}

// One JACK cycle: every client's process callback, sampler channels first, then their passthroughs (graph order).
ZLREF int zlref_process(int nframes, unsigned long long current_usecs, unsigned long long next_usecs, unsigned long long playhead,
                        unsigned long long playhead_usecs, unsigned long long subbeat_usecs) {
    if (!gWorld.ready || nframes > gWorld.nframesMax) return -1;
    gJack.currentUsecs = current_usecs; gJack.nextUsecs = next_usecs; gJack.periodUsecs = (float) (next_usecs - current_usecs);
    zlref_timer_inject(playhead, playhead_usecs, subbeat_usecs);
    for (int b = 0; b < 12; ++b) { _jack_client* c = channelClient(b); if (c && c->cb) c->cb((jack_nframes_t) nframes, c->arg); }
    for (int b = 0; b < 12; ++b) { _jack_client* c = passthroughClient(b); if (c && c->cb) c->cb((jack_nframes_t) nframes, c->arg); }
    for (auto* c : gJack.clients) if (c->name.rfind("AudioLevels-", 0) == 0 && c->cb) c->cb((jack_nframes_t) nframes, c->arg);
    gJack.currentFrames += (jack_nframes_t) nframes;
    zl_ref_shim_now_msecs += 1;                     // the positions model only compares against "a second ago"
    return 0;
}
// which: 0 raw (the SamplerChannel ports), 1 dry, 2 fx1, 3 fx2.  n + 1 floats are available for the raw ports (index n is
// where quirk Q2 puts the last frame).
// the JACK cycle times alone (the scheduler driver of synctimer_ref.cpp runs SyncTimer's process itself)
ZLREF void zlref_set_cycle(unsigned long long current_usecs, unsigned long long next_usecs) {
    gJack.currentUsecs = current_usecs; gJack.nextUsecs = next_usecs; gJack.periodUsecs = (float) (next_usecs - current_usecs);
}
// the sampler side of a cycle without touching the timer: channels, passthroughs, meters
ZLREF int zlref_process_sampler(int nframes) {
    if (!gWorld.ready || nframes > gWorld.nframesMax) return -1;
    for (int b = 0; b < 12; ++b) { _jack_client* c = channelClient(b); if (c && c->cb) c->cb((jack_nframes_t) nframes, c->arg); }
    for (int b = 0; b < 12; ++b) { _jack_client* c = passthroughClient(b); if (c && c->cb) c->cb((jack_nframes_t) nframes, c->arg); }
    for (auto* c : gJack.clients) if (c->name.rfind("AudioLevels-", 0) == 0 && c->cb) c->cb((jack_nframes_t) nframes, c->arg);
    gJack.currentFrames += (jack_nframes_t) nframes;
    zl_ref_shim_now_msecs += 1;
    return 0;
}
ZLREF void* zlref_clip_ptr(int clip) { return (clip >= 0 && clip < (int) gWorld.clips.size()) ? gWorld.clips[(size_t) clip] : nullptr; }
ZLREF int zlref_clip_index(const void* clip) { for (size_t i = 0; i < gWorld.clips.size(); ++i) if (gWorld.clips[i] == clip) return (int) i; return -1; }
ZLREF int zlref_read(int bus, int which, float* left, float* right, int n) {
    static const char* names[4][2] = {{"left_out", "right_out"}, {"dryOutLeft", "dryOutRight"}, {"wetOutFx1Left", "wetOutFx1Right"}, {"wetOutFx2Left", "wetOutFx2Right"}};
    _jack_client* c = which == 0 ? channelClient(bus) : passthroughClient(bus);
    if (!c || which < 0 || which > 3) return -1;
    memcpy(left, portOf(c, names[which][0])->buf.data(), sizeof(float) * (size_t) n);
    memcpy(right, portOf(c, names[which][1])->buf.data(), sizeof(float) * (size_t) n);
    return 0;
}
// ClipAudioSourcePositionsModel: peakGain() (hysteresis and all), firstProgress(), and the 32 slots (id, gain, progress)
ZLREF float zlref_clip_peak_gain(int clip) { return gWorld.clips[(size_t) clip]->playbackPositionsModel()->peakGain(); }
ZLREF double zlref_clip_first_progress(int clip) { return gWorld.clips[(size_t) clip]->playbackPositionsModel()->firstProgress(); }
ZLREF int zlref_clip_positions(int clip, long long* ids, float* gains, float* progress) {
    ClipAudioSourcePositionsModel* m = gWorld.clips[(size_t) clip]->playbackPositionsModel();
    const int n = m->rowCount(QModelIndex());
    for (int i = 0; i < n; ++i) {
        ids[i] = m->data(m->index(i), ClipAudioSourcePositionsModel::PositionIDRole).value<qint64>();
        gains[i] = m->data(m->index(i), ClipAudioSourcePositionsModel::PositionGainRole).value<float>();
        progress[i] = m->data(m->index(i), ClipAudioSourcePositionsModel::PositionProgressRole).value<float>();
    }
    return n;
}
ZLREF void zlref_set_clock_msecs(long long t) { zl_ref_shim_now_msecs = t; }
// jack_connect by full port names; a destination with several sources gets their sum in connection order
ZLREF int zlref_connect(const char* src, const char* dst) {
    _jack_port* s = findPort(src); _jack_port* d = findPort(dst);
    if (!s || !d) return -1;
    if (d->source && d->mix.empty()) d->mix.push_back(d->source);
    if (!d->mix.empty()) d->mix.push_back(s); else d->source = s;
    return 0;
}

}  // extern "C"
