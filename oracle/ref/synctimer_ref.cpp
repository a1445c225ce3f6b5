// oracle/ref/synctimer_ref.cpp — TEST INFRASTRUCTURE.  The reference's own lib/SyncTimer.cpp, compiled in place by including
// it here (its SyncTimerPrivate / SyncTimerThread / StepData are file-local), against the Qt / JACK stand-ins of shim/:
// the 96-ppqn step ring, scheduleClipCommand's same-step merge, queueClipToStart/StopOnChannel, SyncTimerPrivate::process
// (the per-period walk that hands commands to SamplerSynth with the current tick and advances the playhead), setBpm, the
// command pool (row n2 of SURVEY.md section 8f, and the three getters the voice of row a1 reads).
//
// Supplied around it: TransportManager (JACK transport master: two empty methods), the signal bodies moc would generate, and
// what Qt's event loop would do - the timer thread's tick (hiResTimerCallback) and the pool refill are called by the driver
// once per period instead of by threads and queued signals.  Nothing under libzl_b200/ uses this file.
// everything SyncTimer.cpp includes comes in first, with its access specifiers intact (the standard library does not survive
// "#define private public"); only the reference's own classes are opened up
#include <sched.h>
#include <sys/mman.h>
#include <chrono>
#include <iostream>
#include <sstream>
#include <QDebug>
#include <QHash>
#include <QMutex>
#include <QProcess>
#include <QThread>
#include <QTimer>
#include <QWaitCondition>
#include <jack/jack.h>
#include <jack/statistics.h>
#include <jack/midiport.h>
#include "JUCEHeaders.h"
#define private public
#define protected public
#include "SyncTimer.h"
#include "SamplerSynth.h"
#include "TransportManager.h"
#undef private
#undef protected
#include "ClipAudioSource.h"
#include "ClipCommand.h"
#include "libzl.h"
#include "Helper.h"
#include "MidiRouter.h"
#include "TimerCommand.h"
#define private public
#include "SyncTimer.cpp"
#undef private

// TransportManager: the JACK transport side of playback, not on the path
TransportManager::TransportManager(SyncTimer* parent) : QObject(parent) {}
TransportManager::~TransportManager() {}
void TransportManager::initialize() {}
void TransportManager::restartTransport() {}

// signals: where moc would have generated emitters
void SyncTimerThread::timeout() {}
void SyncTimer::bpmChanged() {}
void SyncTimer::scheduleAheadAmountChanged() {}
void SyncTimer::clipCommandSent(ClipCommand*) {}
void SyncTimer::timerCommand(TimerCommand*) {}
void SyncTimer::timerRunningChanged() {}
void SyncTimer::addedHardwareInputDevice(const QString&, const QString&) {}
void SyncTimer::removedHardwareInputDevice(const QString&, const QString&) {}
void SyncTimer::addedHardwareOutputDevice(const QString&, const QString&) {}
void SyncTimer::removedHardwareOutputDevice(const QString&, const QString&) {}
void SyncTimer::pleaseStartPlayback() {}
void SyncTimer::pleaseStopPlayback() {}

// lib/libzl.cpp:40,311: a global that initJuce() fills AFTER SyncTimer::instance() has returned (:388-389) - SamplerSynth's own
// constructor therefore sees nullptr, the voices (made later by SamplerSynth::initialize, :400) see the timer
SyncTimer* syncTimer{nullptr};
extern "C" QObject* SyncTimer_instance() { return syncTimer; }
// SyncTimer's constructor connects pausedChanged to "d->isPaused = timerThread->isPaused()" with the SyncTimer as context; both
// objects live in the thread that calls start()/stop(), so Qt delivers it directly, inside setPaused() - in particular BEFORE
// stop() walks the ring (:884-917).  The emitter does that here.
void SyncTimerThread::pausedChanged() { if (syncTimer) syncTimer->d->isPaused = paused; }
void zlref_timer_boot() { if (!syncTimer) syncTimer = SyncTimer::instance(); }

namespace {
// what the 50 ms objectGarbageHandler does (lib/SyncTimer.cpp:312-335): used commands go back into the pool
void refillPools(SyncTimerPrivate* d) {
    QMutableListIterator<ClipCommand*> it(d->freshClipCommands);
    while (it.hasNext() && d->clipCommandsToDelete.count() > 0) {
        ClipCommand* value = it.next();
        if (value == nullptr) {
            ClipCommand* refreshed = d->clipCommandsToDelete.takeFirst();
            ClipCommand::clear(refreshed);
            it.setValue(refreshed);
        }
    }
    QMutableListIterator<TimerCommand*> jt(d->freshTimerCommands);
    while (jt.hasNext() && d->timerCommandsToDelete.count() > 0) {
        TimerCommand* value = jt.next();
        if (value == nullptr) {
            TimerCommand* refreshed = d->timerCommandsToDelete.takeFirst();
            TimerCommand::clear(refreshed);
            jt.setValue(refreshed);
        }
    }
}
}  // namespace

// hooks used by libzl_ref.cpp (the voice scenarios set the playhead values the voice reads directly)
void zlref_timer_inject(unsigned long long playhead, unsigned long long playhead_usecs, unsigned long long subbeat_usecs) {
    SyncTimerPrivate* d = SyncTimer::instance()->d;
    timerThread->paused = false;                    // the getters report jackPlayhead / jackNextPlaybackPosition when running (:990-1004)
    d->isPaused = false;
    d->jackPlayhead = playhead; d->jackNextPlaybackPosition = playhead_usecs; d->jackSubbeatLengthInMicroseconds = subbeat_usecs;
    refillPools(d);
}
void zlref_timer_handle(ClipCommand* c, unsigned long long tick) { SamplerSynth::instance()->handleClipCommand(c, tick); }   // SyncTimerPrivate is SamplerSynth's friend
void zlref_timer_enable(int channel, bool enabled) { SamplerSynth::instance()->setChannelEnabled(channel, enabled); }

extern "C" {
void zlref_set_cycle(unsigned long long current_usecs, unsigned long long next_usecs);
int zlref_process_sampler(int nframes);
void* zlref_clip_ptr(int clip);
int zlref_clip_index(const void* clip);
#define ZLREF __attribute__((visibility("default")))

// ---- the scheduler driven for real (row n2) ----
ZLREF int zlref_sched_start(int bpm) {              // SyncTimer::start (resume() -> pausedChanged -> isPaused, see above)
    SyncTimer* t = SyncTimer::instance();
    t->start(bpm);
    return 0;
}
ZLREF int zlref_sched_stop(void) {
    SyncTimer* t = SyncTimer::instance();
    t->stop();
    return 0;
}
ZLREF int zlref_sched_set_bpm(int bpm) { SyncTimer::instance()->setBpm((quint64) bpm); return 0; }
ZLREF int zlref_sched_schedule(int clip, int midi_note, int midi_channel, int start_playback, int stop_playback, int change_slice, int slice,
                               int change_looping, int looping, int change_volume, float volume, unsigned long long delay) {
    SyncTimer* t = SyncTimer::instance();
    ClipCommand* c = t->getClipCommand();
    if (!c) return -1;
    c->clip = static_cast<ClipAudioSource*>(zlref_clip_ptr(clip)); c->midiNote = midi_note; c->midiChannel = midi_channel;
    c->startPlayback = start_playback; c->stopPlayback = stop_playback; c->changeSlice = change_slice; c->slice = slice;
    c->changeLooping = change_looping; c->looping = looping; c->changeVolume = change_volume; c->volume = volume;
    t->scheduleClipCommand(c, delay);
    return 0;
}
ZLREF int zlref_sched_queue_start(int clip, int midi_channel) { SyncTimer::instance()->queueClipToStartOnChannel(static_cast<ClipAudioSource*>(zlref_clip_ptr(clip)), midi_channel); return 0; }
ZLREF int zlref_sched_queue_stop(int clip, int midi_channel) { SyncTimer::instance()->queueClipToStopOnChannel(static_cast<ClipAudioSource*>(zlref_clip_ptr(clip)), midi_channel); return 0; }
ZLREF int zlref_sched_channel_enabled(unsigned long long delay, int channel, int enabled) {
    SyncTimer::instance()->scheduleTimerCommand(delay, TimerCommand::SamplerChannelEnabledStateOperation, channel, enabled);
    return 0;
}
// One JACK cycle with the real timer in front: SyncTimerPrivate::process, then the sampler clients, then what the timer
// thread and the event loop would have done by the next cycle.  out[3] = playhead, playhead_usecs, subbeat_usecs as the voices
// of this cycle saw them; the commands handed over in this cycle are copied to cmds (11 ints + volume each, up to cap).
ZLREF int zlref_sched_cycle(int nframes, unsigned long long current_usecs, unsigned long long next_usecs, unsigned long long* out,
                            int* cmd_ints, float* cmd_volumes, int cap) {
    SyncTimer* t = SyncTimer::instance();
    SyncTimerPrivate* d = t->d;
    zlref_set_cycle(current_usecs, next_usecs);
    ClipCommandRingEntry* before = d->sentOutClipsWriteHead;
    d->process((jack_nframes_t) nframes);
    int n = 0;
    for (ClipCommandRingEntry* e = before; e != d->sentOutClipsWriteHead; e = e->next) {
        const ClipCommand* c = e->clipCommand;
        if (n < cap && c) {
            int* o = cmd_ints + 11 * n;
            o[0] = zlref_clip_index(c->clip); o[1] = c->midiNote; o[2] = c->midiChannel; o[3] = c->startPlayback; o[4] = c->stopPlayback;
            o[5] = c->changeSlice; o[6] = c->slice; o[7] = c->changeLooping; o[8] = c->looping; o[9] = c->changeVolume; o[10] = 0;
            cmd_volumes[n] = c->volume;
        }
        ++n;
    }
    out[0] = t->jackPlayhead(); out[1] = t->jackPlayheadUsecs(); out[2] = t->jackSubbeatLengthInMicroseconds();
    if (zlref_process_sampler(nframes) != 0) return -1;
    if (!d->isPaused) d->hiResTimerCallback();      // the timer thread: at least one tick per cycle
    else while (d->sentOutClipsReadHead->clipCommand) { d->sentOutClipsReadHead->clipCommand = nullptr; d->sentOutClipsReadHead = d->sentOutClipsReadHead->next; }
    refillPools(d);
    return n;
}
ZLREF int zlref_sched_state(unsigned long long* out /* cumulativeBeat, beat, scheduleAhead, bpm */) {
    SyncTimer* t = SyncTimer::instance();
    out[0] = t->cumulativeBeat(); out[1] = (unsigned long long) t->beat(); out[2] = t->scheduleAheadAmount(); out[3] = t->getBpm();
    return 0;
}
}
