// oracle/ref/audiolevels_ref.cpp — TEST INFRASTRUCTURE.  The reference's own lib/AudioLevels.cpp, compiled in place by
// including it here (so that its file-local AudioLevelsChannel / AudioLevelsPrivate are visible to the driver below), against
// the Qt / JACK stand-ins of shim/ (row a10: AudioLevelsChannel::process, AudioLevels::timerCallback - the integer peak scan,
// the 10000-per-tick decay, convertTodbFS, the playback hold, add()).  Nothing under libzl_b200/ uses this file.
#define private public
#include "AudioLevels.h"
#undef private
#include "AudioLevels.cpp"

#include <filesystem>

bool QDir::exists(const QString& path) const { return std::filesystem::exists(path.toStdString()); }
bool QDir::mkpath(const QString& path) const { std::error_code ec; std::filesystem::create_directories(path.toStdString(), ec); return !ec; }

// signals: where moc would have generated emitters
void AudioLevels::audioLevelsChanged() {}
void AudioLevels::recordGlobalPlaybackChanged() {}
void AudioLevels::channelsToRecordChanged() {}
void AudioLevels::shouldRecordPortsChanged() {}
void AudioLevels::isRecordingChanged() {}

extern "C" {
// creates the 13 AudioLevels JACK clients (lib/AudioLevels.cpp:282-330); their ports are wired by zlref_connect
__attribute__((visibility("default"))) int zlref_levels_init(void) { return AudioLevels::instance() != nullptr ? 0 : -1; }
// AudioLevels::timerCallback, then for each of the 13 channels: peakA, peakB (ints) and out[13][6] =
// dB A, dB B, combined (playback / track levels), hold A (linear), hold B (linear), 0
__attribute__((visibility("default"))) int zlref_levels_tick(int* peaks /* [13][2] */, float* out /* [13][6] */) {
    AudioLevels* a = AudioLevels::instance();
    a->timerCallback();
    for (int i = 0; i < 13; ++i) {
        AudioLevelsChannel* c = a->d->audioLevelsChannels[i];
        peaks[2 * i] = c->peakA; peaks[2 * i + 1] = c->peakB;
        float* o = out + 6 * i;
        o[3] = c->peakAHoldSignal; o[4] = c->peakBHoldSignal; o[5] = 0.0f;
        if (i == 0) { o[0] = a->captureA; o[1] = a->captureB; o[2] = 0.0f; }
        else if (i == 1) { o[0] = a->playbackA; o[1] = a->playbackB; o[2] = a->playback; }
        else if (i == 2) { o[0] = a->recordingA; o[1] = a->recordingB; o[2] = 0.0f; }
        else { o[0] = a->channelsA[i - 3]; o[1] = a->channelsB[i - 3]; o[2] = a->d->levels[i - 3].toFloat(); }
    }
    return 0;
}
__attribute__((visibility("default"))) float zlref_levels_add(float a, float b) { return AudioLevels::instance()->add(a, b); }
}
