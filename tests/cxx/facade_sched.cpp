// Row n2 through the reference-named classes: a session driven by SyncTimer (the facade over zl_sched_*) instead of a
// hand-written command stream - commands scheduled while paused and while running, clips queued to the next bar, a tempo
// change, a channel switched off by a timer command, a queued stop, stop() - period by period (process) and in batches
// (render).  Writes every period's clock and the raw ports of five channels; tests/test_facade.py replays the same calls
// on the oracle through the C ABI and compares the bytes.
#include "zl/SyncTimer.h"
#include "zl/SamplerSynth.h"
#include "zl/ClipAudioSource.h"

#include <cstdio>
#include <cstring>
#include <vector>

static FILE* out = nullptr;
static SamplerSynth* synth = nullptr;
static SyncTimer* timer = nullptr;
static const int kFrames = 256;
static const uint64_t kPeriod = 5333;                  // llround(256e6 / 48000)
static uint64_t periodNo = 0;

static bool dump(int nblocks) {
    const size_t n = (size_t) synth->framesInLastRender();
    if (n != (size_t) nblocks * kFrames) return false;
    std::vector<float> l(n), r(n);
    const int channels[5] = {-2, 0, 1, 2, 3};
    for (int ch : channels) {
        if (!synth->readChannel(ch, ZL_BUS_RAW, l.data(), r.data())) return false;
        fwrite(l.data(), 4, n, out); fwrite(r.data(), 4, n, out);
    }
    const uint64_t st[4] = {timer->jackPlayhead(), timer->jackPlayheadUsecs(), timer->jackSubbeatLengthInMicroseconds(), timer->cumulativeBeat()};
    fwrite(st, 8, 4, out);
    return true;
}
static bool one_by_one(int periods) {
    for (int i = 0; i < periods; ++i, ++periodNo) {
        const uint64_t cur = 1000000 + periodNo * kPeriod;
        if (!timer->process(kFrames, cur, cur + kPeriod) || !dump(1)) return false;
    }
    return true;
}
static bool batched(int periods, int batch) {
    for (int done = 0; done < periods; done += batch) {
        const int nb = periods - done < batch ? periods - done : batch;
        if (!timer->render(nb, kFrames, 1000000 + periodNo * kPeriod) || !dump(nb)) return false;
        periodNo += (uint64_t) nb;
    }
    return true;
}
static ClipCommand* play(ClipAudioSource* clip, int channel, int note, float volume, bool loop) {
    ClipCommand* c = timer->getClipCommand();
    c->clip = clip; c->midiChannel = channel; c->midiNote = note; c->changeVolume = true; c->volume = volume;
    c->looping = loop; c->stopPlayback = loop; c->startPlayback = true;
    return c;
}

int main(int argc, char** argv) {
    if (argc < 4) return 1;
    synth = SamplerSynth::instance();
    if (!synth->initialize(nullptr, 48000.0, kFrames, 16)) {
        fprintf(stderr, "SamplerSynth::initialize: %s\n", synth->lastError());
        return strstr(synth->lastError(), "no CUDA device") ? 3 : 1;
    }
    timer = SyncTimer::instance();
    ClipAudioSource clip(nullptr, nullptr, argv[1]);
    ClipAudioSource loop(nullptr, nullptr, argv[2]);
    if (clip.deviceClip() < 0 || loop.deviceClip() < 0) return 1;
    loop.setADSRParameters(ADSRParameters(0.002f, 0.05f, 0.7f, 0.03f));
    loop.setLength(1.5f, 120);
    loop.setVolumeAbsolute(0.8f);
    if (timer->getMultiplier() != 96 || timer->getInterval(120) != 5 || timer->secondsToSubbeatCount(120, 0.5f) != 96 || timer->timerRunning()) return 1;
    out = fopen(argv[3], "wb");
    if (!out) return 1;
    uint64_t seen = 0;
    timer->clipCommandSent = [&](ClipCommand* c) { if (c->clip == &clip || c->clip == &loop) ++seen; };

    // paused: delays count from the read head
    timer->scheduleClipCommand(play(&clip, -2, 60, 0.7f, false), 0);
    timer->scheduleClipCommand(play(&loop, 0, 62, 0.5f, true), 3);
    { ClipCommand* c = timer->getClipCommand(); c->clip = &loop; c->midiChannel = 0; c->midiNote = 62; c->changeVolume = true; c->volume = 0.9f;
      timer->scheduleClipCommand(c, 3); }                            // merges into the one above
    timer->queueClipToStartOnChannel(&loop, 1);
    if (!one_by_one(12)) return 1;
    timer->start(120);
    if (!timer->timerRunning() || timer->getBpm() != 120) return 1;
    timer->scheduleClipCommand(play(&clip, 2, 64, 0.6f, false), 10);
    timer->queueClipToStartOnChannel(&clip, 3);
    if (!batched(40, 8)) return 1;
    timer->setBpm(90);
    timer->scheduleTimerCommand(4, 8, 3, 0);                          // TimerCommand::SamplerChannelEnabledStateOperation: channel 3 off
    timer->queueClipToStopOnChannel(&loop, 1);
    timer->scheduleClipCommand(play(&loop, 2, 67, 0.4f, false), 96);
    if (!one_by_one(30)) return 1;
    timer->scheduleClipCommand(play(&clip, 0, 55, 0.9f, false), 400);  // still waiting when the timer stops: comes out muted
    timer->stop();
    if (!batched(10, 10)) return 1;
    const uint64_t tail[3] = {timer->commandsSent(), seen, timer->periodsProcessed()};
    fwrite(tail, 8, 3, out);
    fclose(out);
    return tail[0] == tail[1] ? 0 : 1;
}
