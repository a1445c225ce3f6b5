// libzl_b200/host/libzl_capi.cpp — the subset of the reference's own C ABI (lib/libzl.h:18-179, implemented in lib/libzl.cpp)
// that concerns the sampler render path, with IDENTICAL names and signatures, so a ctypes host written for libzl.so
// (test/playtest.py:25-47) binds unchanged.  Thin forwards to the façade classes of include/zl/libzl_facade.h; file loading
// is WAV only (the library's own reader), everything tracktion time-stretch related is a stored-but-ignored parameter.
#include "../../include/zl/libzl_facade.h"

#include <cmath>
#include <vector>

namespace { std::vector<ClipAudioSource*> gClips; }

extern "C" {
#define ZLX __attribute__((visibility("default")))

ZLX void initJuce() { SamplerSynth::instance()->initialize(nullptr); }                     // lib/libzl.cpp:358-410
ZLX void shutdownJuce() {}
ZLX ClipAudioSource* ClipAudioSource_new(const char* filepath, bool muted) {               // :107-128 (ids from 1)
    ClipAudioSource* c = new ClipAudioSource(nullptr, nullptr, filepath, muted);
    gClips.push_back(c);
    c->setId((int) gClips.size());
    return c;
}
ZLX ClipAudioSource* ClipAudioSource_byID(int id) { return (id > 0 && id <= (int) gClips.size()) ? gClips[(size_t) id - 1] : nullptr; }
ZLX void ClipAudioSource_destroy(ClipAudioSource* c) {                                      // :258-267
    for (auto& p : gClips) if (p == c) p = nullptr;
    delete c;
}
ZLX int ClipAudioSource_id(ClipAudioSource* c) { return c->id(); }
ZLX void ClipAudioSource_setProgressCallback(ClipAudioSource* c, void (*f)(float)) { c->setProgressCallback(f); }
ZLX void ClipAudioSource_play(ClipAudioSource* c, bool loop) { c->play(loop); }
ZLX void ClipAudioSource_stop(ClipAudioSource* c) { c->stop(); }
ZLX void ClipAudioSource_playOnChannel(ClipAudioSource* c, bool loop, int midiChannel) { c->play(loop, midiChannel); }
ZLX void ClipAudioSource_stopOnChannel(ClipAudioSource* c, int midiChannel) { c->stop(midiChannel); }
ZLX float ClipAudioSource_getDuration(ClipAudioSource* c) { return c->getDuration(); }
ZLX const char* ClipAudioSource_getFileName(ClipAudioSource* c) { return c->getFileName(); }
ZLX void ClipAudioSource_setStartPosition(ClipAudioSource* c, float s) { c->setStartPosition(s); }
ZLX void ClipAudioSource_setLength(ClipAudioSource* c, float beat, int bpm) { c->setLength(beat, bpm); }
ZLX void ClipAudioSource_setPan(ClipAudioSource* c, float pan) { c->setPan(pan); }
ZLX void ClipAudioSource_setSpeedRatio(ClipAudioSource* c, float r) { c->setSpeedRatio(r); }
ZLX void ClipAudioSource_setPitch(ClipAudioSource* c, float p) { c->setPitch(p); }
ZLX void ClipAudioSource_setGain(ClipAudioSource* c, float db) { c->setGain(db); }
ZLX void ClipAudioSource_setVolume(ClipAudioSource* c, float vol) { c->setVolume(vol); }
ZLX void ClipAudioSource_setAudioLevelChangedCallback(ClipAudioSource* c, void (*f)(float)) { c->setAudioLevelChangedCallback(f); }
ZLX void ClipAudioSource_setSlices(ClipAudioSource* c, int slices) { c->setSlices(slices); }
ZLX int ClipAudioSource_keyZoneStart(ClipAudioSource* c) { return c->keyZoneStart(); }
ZLX void ClipAudioSource_setKeyZoneStart(ClipAudioSource* c, int v) { c->setKeyZoneStart(v); }
ZLX int ClipAudioSource_keyZoneEnd(ClipAudioSource* c) { return c->keyZoneEnd(); }
ZLX void ClipAudioSource_setKeyZoneEnd(ClipAudioSource* c, int v) { c->setKeyZoneEnd(v); }
ZLX int ClipAudioSource_rootNote(ClipAudioSource* c) { return c->rootNote(); }
ZLX void ClipAudioSource_setRootNote(ClipAudioSource* c, int v) { c->setRootNote(v); }
ZLX float ClipAudioSource_adsrAttack(ClipAudioSource* c) { return c->adsrAttack(); }
ZLX void ClipAudioSource_setADSRAttack(ClipAudioSource* c, float v) { c->setADSRAttack(v); }
ZLX float ClipAudioSource_adsrDecay(ClipAudioSource* c) { return c->adsrDecay(); }
ZLX void ClipAudioSource_setADSRDecay(ClipAudioSource* c, float v) { c->setADSRDecay(v); }
ZLX float ClipAudioSource_adsrSustain(ClipAudioSource* c) { return c->adsrSustain(); }
ZLX void ClipAudioSource_setADSRSustain(ClipAudioSource* c, float v) { c->setADSRSustain(v); }
ZLX float ClipAudioSource_adsrRelease(ClipAudioSource* c) { return c->adsrRelease(); }
ZLX void ClipAudioSource_setADSRRelease(ClipAudioSource* c, float v) { c->setADSRRelease(v); }
ZLX void stopClips(int size, ClipAudioSource** clips) { for (int i = 0; i < size; ++i) if (clips[i]) clips[i]->stop(); }   // :412-424
ZLX float dBFromVolume(float vol) { return vol > 0.0f ? 20.0f * std::log(vol) + 6.0f : -100.0f; }   // te::volumeFaderPositionToDB  :426-429

// JackPassthrough_*(channel: -1 = global, 0..9 tracks)   lib/libzl.cpp:476-575
static JackPassthrough pt(int channel) { return JackPassthrough(channel < 0 ? -2 : channel); }
ZLX float JackPassthrough_getPanAmount(int channel) { return pt(channel).panAmount(); }
ZLX float JackPassthrough_getWetFx1Amount(int channel) { return pt(channel).wetFx1Amount(); }
ZLX float JackPassthrough_getWetFx2Amount(int channel) { return pt(channel).wetFx2Amount(); }
ZLX float JackPassthrough_getDryAmount(int channel) { return pt(channel).dryAmount(); }
ZLX float JackPassthrough_getMuted(int channel) { return pt(channel).muted() ? 1.0f : 0.0f; }
ZLX void JackPassthrough_setPanAmount(int channel, float v) { pt(channel).setPanAmount(v); }
ZLX void JackPassthrough_setWetFx1Amount(int channel, float v) { pt(channel).setWetFx1Amount(v); }
ZLX void JackPassthrough_setWetFx2Amount(int channel, float v) { pt(channel).setWetFx2Amount(v); }
ZLX void JackPassthrough_setDryAmount(int channel, float v) { pt(channel).setDryAmount(v); }
ZLX void JackPassthrough_setMuted(int channel, bool v) { pt(channel).setMuted(v); }
}
