// zl_sched.cpp — row n2 of SURVEY.md section 8f: the command-scheduling front-end, i.e. the CALLER of the render path.
//
// The reference's SyncTimer keeps a ring of 32768 steps at 96 steps per quarter note; clip commands are scheduled some steps
// ahead (SyncTimer::scheduleClipCommand merges a command into an equivalent one already waiting on that step,
// lib/SyncTimer.cpp:1011-1048), and once per JACK period SyncTimerPrivate::process (:452-702) walks the steps whose time has
// come, hands their commands to SamplerSynth::handleClipCommand together with the current tick (:553-558) and advances the
// playhead the voices read (jackPlayhead / jackPlayheadUsecs / jackSubbeatLengthInMicroseconds, :990-1009).
//
// Here the same state machine runs on the host as plain C++ behind the C ABI (zl_sched_*): zl_sched_advance takes N periods
// at once and returns (a) the zl_clip_commands that fall due, each with the period it belongs to and the tick the reference
// would stamp it with, ready for zl_command_batch, and (b) the zl_block_clock of every period, ready for zl_render - so a
// whole sketch is rendered offline without a hand-written command stream.  No device work happens here (the survey marks
// this row as the caller of the path); the arithmetic that matters is the reference's mix of integer and floating-point
// microseconds, kept operation for operation so that block clocks and ticks agree exactly.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <new>
#include <vector>

#include "../../include/zl_b200.h"

namespace {

constexpr uint64_t kNanosecondsPerMinute = 60000000000ull;     // lib/SyncTimer.cpp:92
constexpr uint64_t kBeatSubdivisions = 96;                      // :95
constexpr uint64_t kStepRingCount = 32768;                      // :276
constexpr uint64_t kBpmMinimum = 50, kBpmMaximum = 200;         // :28-29

inline uint64_t subbeatCountToNanoseconds(uint64_t bpm, uint64_t subBeatCount) { return (subBeatCount * kNanosecondsPerMinute) / (bpm * kBeatSubdivisions); }   // :180-183
inline float nanosecondsToSubbeatCount(uint64_t bpm, uint64_t nanoseconds) { return (float) (nanoseconds / (kNanosecondsPerMinute / (bpm * kBeatSubdivisions))); }   // :184-187

enum { kOpSetBpm = 1, kOpChannelEnabled = 2 };                  // the TimerCommand operations that concern the path (:600-640)
struct TimerCmd { int op, p1, p2; };

struct Step {                                                   // StepData  :42-77
    std::vector<zl_clip_command> clipCommands;
    std::vector<TimerCmd> timerCommands;
    bool played = true;
    void ensureFresh() { if (played) { played = false; timerCommands.clear(); clipCommands.clear(); } }
};

// ClipCommand::equivalentTo   lib/ClipCommand.h:33-39
inline bool equivalent(const zl_clip_command& a, const zl_clip_command& b) {
    return a.clip == b.clip
        && ((a.change_slice && b.change_slice && a.slice == b.slice)
            || (!a.change_slice && !b.change_slice && a.midi_note == b.midi_note && a.midi_channel == b.midi_channel));
}

}  // namespace

struct zl_sched {
    std::vector<Step> ring = std::vector<Step>(kStepRingCount);
    uint64_t stepReadHead = 0;                                  // index of the next step to be read
    uint64_t stepNextPlaybackPosition = 0;
    uint64_t jackPlayhead = 0, jackNextPlaybackPosition = 0, jackSubbeatLengthInMicroseconds = 0, jackLatency = 0;
    double jackPlayheadBpm = 120;
    uint64_t stepReadHeadOnStart = 0, cumulativeBeat = 0, scheduleAheadAmount = 0, bpm = 120;
    int beat = 0;
    bool isPaused = true;
    double sampleRate = 48000.0;

    void updateScheduleAheadAmount() { scheduleAheadAmount = (uint64_t) (nanosecondsToSubbeatCount(bpm, (uint64_t) (jackLatency * (float) 1000000)) + 1); }   // :712-715

    Step* delayedStep(uint64_t delay, bool ensureFresh = true) {           // :346-362
        uint64_t step;
        if (isPaused) step = (stepReadHead + delay + 1) % kStepRingCount;
        else step = (stepReadHeadOnStart + std::max(cumulativeBeat + delay, jackPlayhead + 1)) % kStepRingCount;
        Step* s = &ring[step];
        if (ensureFresh) s->ensureFresh();
        return s;
    }

    void scheduleClipCommand(const zl_clip_command& command, uint64_t delay) {   // :1011-1048
        Step* stepData = delayedStep(delay);
        bool foundExisting = false;
        for (zl_clip_command& existing : stepData->clipCommands) {
            if (equivalent(existing, command)) {
                if (command.change_looping) { existing.looping = command.looping; existing.change_looping = 1; }
                if (command.change_pitch) { existing.pitch_change = command.pitch_change; existing.change_pitch = 1; }
                if (command.change_speed) { existing.speed_ratio = command.speed_ratio; existing.change_speed = 1; }
                if (command.change_gain_db) { existing.gain_db = command.gain_db; existing.change_gain_db = 1; }
                if (command.change_volume) { existing.volume = command.volume; existing.change_volume = 1; }
                if (command.start_playback) existing.start_playback = 1;
                foundExisting = true;
            }
        }
        if (!foundExisting) stepData->clipCommands.push_back(command);
    }

    void setBpm(uint64_t newBpm) {                                          // :955-975
        if (bpm != newBpm) {
            bpm = newBpm;
            jackSubbeatLengthInMicroseconds = subbeatCountToNanoseconds(bpm, 1) / 1000;
            updateScheduleAheadAmount();
            delayedStep(0)->timerCommands.push_back(TimerCmd{kOpSetBpm, (int) newBpm, 0});
        }
    }

    void timerTick() {                                                      // hiResTimerCallback  :392-415
        while (cumulativeBeat < (jackPlayhead + (scheduleAheadAmount * 2))) {
            beat = (beat + 1) % (int) (kBeatSubdivisions * 4);
            ++cumulativeBeat;
        }
    }

    // SyncTimerPrivate::process for one period  :452-702.  Events go to `out`; returns false when it overflows.
    bool process(uint32_t nframes, uint64_t current_usecs, uint64_t next_usecs, int block, std::vector<zl_sched_event>& out) {
        const float period_usecs = (float) (next_usecs - current_usecs);    // jack_get_cycle_times' period
        (void) period_usecs;
        double thisStepBpm = jackPlayheadBpm;
        double thisStepSubbeatLengthInMicroseconds = double(subbeatCountToNanoseconds((uint64_t) jackPlayheadBpm, 1)) / 1000.0;   // :474
        if (!isPaused) {
            if (jackPlayhead == 0) jackNextPlaybackPosition = current_usecs;   // :481-489
        }
        if (stepNextPlaybackPosition == 0) stepNextPlaybackPosition = current_usecs;   // :491-493
        uint32_t firstAvailableFrame = 0;
        const uint64_t microsecondsPerFrame = (next_usecs - current_usecs) / nframes;
        while (stepNextPlaybackPosition < next_usecs && firstAvailableFrame < nframes) {   // :503
            Step* stepData = &ring[stepReadHead];
            stepReadHead = (stepReadHead + 1) % kStepRingCount;
            if (stepNextPlaybackPosition <= current_usecs) { ++firstAvailableFrame; }                  // :508-514
            else {
                const uint64_t rel = microsecondsPerFrame ? (stepNextPlaybackPosition - current_usecs) / microsecondsPerFrame : 0;
                firstAvailableFrame = (uint32_t) std::min<uint64_t>(std::max<uint64_t>(rel, firstAvailableFrame), nframes - 1);
            }
            if (!stepData->played) {
                for (const zl_clip_command& c : stepData->clipCommands) {   // :553-558: handleClipCommand(command, jackPlayhead)
                    zl_sched_event ev;
                    memset(&ev, 0, sizeof ev);
                    ev.kind = ZL_SCHED_CLIP_COMMAND; ev.block = block; ev.tick = jackPlayhead; ev.cmd = c;
                    out.push_back(ev);
                }
                for (const TimerCmd& t : stepData->timerCommands) {         // :563-640, the operations that reach the sampler
                    if (t.op == kOpChannelEnabled) {
                        zl_sched_event ev;
                        memset(&ev, 0, sizeof ev);
                        ev.kind = ZL_SCHED_CHANNEL_ENABLED; ev.block = block; ev.tick = jackPlayhead; ev.channel = t.p1; ev.enabled = t.p2;
                        out.push_back(ev);
                    } else if (t.op == kOpSetBpm) {
                        const uint64_t newBpm = std::min<uint64_t>(std::max<uint64_t>((uint64_t) t.p1, 50), 200);
                        setBpm(newBpm);
                        thisStepBpm = (double) newBpm;
                    }
                }
                stepData->played = true;
            }
            if (jackPlayheadBpm != thisStepBpm) {                           // :645-650: from here on the integer microseconds
                jackPlayheadBpm = thisStepBpm;
                thisStepSubbeatLengthInMicroseconds = (double) (subbeatCountToNanoseconds((uint64_t) jackPlayheadBpm, 1) / 1000);
            }
            if (!isPaused) {                                                // :668-675
                ++jackPlayhead;
                jackNextPlaybackPosition = (uint64_t) ((double) jackNextPlaybackPosition + thisStepSubbeatLengthInMicroseconds);
            }
            stepNextPlaybackPosition = (uint64_t) ((double) stepNextPlaybackPosition + thisStepSubbeatLengthInMicroseconds);   // :677
        }
        return true;
    }
};

extern "C" {

int zl_sched_create(double sample_rate, int32_t buffer_size, zl_sched** out) {
    if (!out || !(sample_rate > 0) || buffer_size < 1) return ZL_ERR_INVALID;
    zl_sched* s = new (std::nothrow) zl_sched;
    if (!s) return ZL_ERR_NOMEM;
    s->sampleRate = sample_rate;
    s->jackSubbeatLengthInMicroseconds = subbeatCountToNanoseconds(s->bpm, 1) / 1000;      // :749
    s->jackLatency = (uint64_t) ((1000 * (double) buffer_size) / sample_rate);                // :769
    s->updateScheduleAheadAmount();
    *out = s;
    return ZL_OK;
}
int zl_sched_destroy(zl_sched* s) { if (!s) return ZL_ERR_INVALID; delete s; return ZL_OK; }

int zl_sched_start(zl_sched* s, int32_t bpm) {                                  // SyncTimer::start  :870-879
    if (!s || bpm < 1) return ZL_ERR_INVALID;
    s->setBpm((uint64_t) bpm);
    s->stepReadHeadOnStart = s->stepReadHead;
    s->isPaused = false;
    return ZL_OK;
}
int zl_sched_stop(zl_sched* s) {                                                // SyncTimer::stop  :881-924
    if (!s) return ZL_ERR_INVALID;
    s->isPaused = true;
    s->beat = 0; s->cumulativeBeat = 0; s->jackPlayhead = 0;
    // The reference walks the whole ring from the read head and re-schedules every waiting command at delay 0 with its volume
    // forced to 0; paused (which it already is at that point) delay 0 is the step behind the read head.  That step collects the
    // read head's commands, is visited second (its commands merge into themselves) and marked played, is cleared by the first
    // command that arrives after that, and ends the walk holding everything that was waiting further out - muted, unplayed,
    // emitted when the read head next moves.  Same walk here, same order.
    for (uint64_t k = 0; k < kStepRingCount; ++k) {
        Step& sd = s->ring[(k + s->stepReadHead) % kStepRingCount];
        if (sd.played) continue;
        for (size_t i = 0; i < sd.clipCommands.size(); ++i) {
            sd.clipCommands[i].change_volume = 1;
            sd.clipCommands[i].volume = 0.0f;
            const zl_clip_command c = sd.clipCommands[i];
            s->scheduleClipCommand(c, 0);
        }
        sd.played = true;
    }
    return ZL_OK;
}
int zl_sched_set_bpm(zl_sched* s, int32_t bpm) { if (!s || bpm < 1) return ZL_ERR_INVALID; s->setBpm((uint64_t) bpm); return ZL_OK; }

int zl_sched_schedule(zl_sched* s, const zl_clip_command* cmd, uint64_t delay) {
    if (!s || !cmd) return ZL_ERR_INVALID;
    s->scheduleClipCommand(*cmd, delay);
    return ZL_OK;
}
int zl_sched_schedule_channel_enabled(zl_sched* s, uint64_t delay, int32_t channel, int32_t enabled) {   // TimerCommand::SamplerChannelEnabledStateOperation
    if (!s) return ZL_ERR_INVALID;
    s->delayedStep(delay)->timerCommands.push_back(TimerCmd{kOpChannelEnabled, channel, enabled});
    return ZL_OK;
}
int zl_sched_queue_start(zl_sched* s, int32_t clip, int32_t midi_channel) {    // SyncTimer::queueClipToStartOnChannel  :816-832
    if (!s) return ZL_ERR_INVALID;
    zl_clip_command c;
    zl_clip_command_default(&c);
    c.clip = clip; c.midi_channel = midi_channel; c.midi_note = 60; c.change_volume = 1; c.volume = 1.0f; c.looping = 1;
    c.stop_playback = 1; c.start_playback = 1;
    const uint64_t bar = kBeatSubdivisions * 4;
    const uint64_t nextZeroBeat = s->isPaused ? 0 : bar - (s->cumulativeBeat % bar);
    s->scheduleClipCommand(c, s->cumulativeBeat + nextZeroBeat < s->jackPlayhead ? nextZeroBeat + bar : nextZeroBeat);
    return ZL_OK;
}
int zl_sched_queue_stop(zl_sched* s, int32_t clip, int32_t midi_channel) {     // SyncTimer::queueClipToStopOnChannel  :834-860
    if (!s) return ZL_ERR_INVALID;
    for (Step& sd : s->ring)                                                     // the first waiting command of the clip on every step goes
        if (!sd.played)
            for (size_t i = 0; i < sd.clipCommands.size(); ++i)
                if (sd.clipCommands[i].clip == clip) { sd.clipCommands.erase(sd.clipCommands.begin() + (long) i); break; }
    zl_clip_command c;
    zl_clip_command_default(&c);
    c.clip = clip; c.midi_channel = midi_channel; c.midi_note = 60; c.stop_playback = 1;
    s->delayedStep(0)->clipCommands.push_back(c);                                // no merge: appended as it is
    return ZL_OK;
}

int zl_sched_advance(zl_sched* s, int32_t nblocks, int32_t nframes, zl_block_clock* clocks, zl_sched_event* events, int32_t cap, int32_t* count) {
    if (!s || nblocks < 1 || nframes < 1 || !clocks || !count || cap < 0 || (cap && !events)) return ZL_ERR_INVALID;
    std::vector<zl_sched_event> out;
    for (int b = 0; b < nblocks; ++b) {
        s->process((uint32_t) nframes, clocks[b].current_usecs, clocks[b].next_usecs, b, out);
        // what the voices of this period read  (lib/SyncTimer.cpp:990-1009, Q4: other quantities while paused)
        clocks[b].playhead = s->isPaused ? s->stepReadHead : s->jackPlayhead;
        clocks[b].playhead_usecs = s->isPaused ? s->stepNextPlaybackPosition : s->jackNextPlaybackPosition;
        clocks[b].subbeat_usecs = s->jackSubbeatLengthInMicroseconds;
        if (!s->isPaused) s->timerTick();                                        // the timer thread runs at least once per period
    }
    *count = (int32_t) out.size();
    if ((int32_t) out.size() > cap) return ZL_ERR_STATE;                          // *count says how much room is needed; the ring has moved on
    if (!out.empty()) memcpy(events, out.data(), out.size() * sizeof(zl_sched_event));
    return ZL_OK;
}

int zl_sched_state(zl_sched* s, uint64_t* playhead, uint64_t* playhead_usecs, uint64_t* subbeat_usecs, uint64_t* cumulative_beat,
                   int32_t* beat, uint64_t* schedule_ahead) {
    if (!s) return ZL_ERR_INVALID;
    if (playhead) *playhead = s->isPaused ? s->stepReadHead : s->jackPlayhead;
    if (playhead_usecs) *playhead_usecs = s->isPaused ? s->stepNextPlaybackPosition : s->jackNextPlaybackPosition;
    if (subbeat_usecs) *subbeat_usecs = s->jackSubbeatLengthInMicroseconds;
    if (cumulative_beat) *cumulative_beat = s->cumulativeBeat;
    if (beat) *beat = s->beat;
    if (schedule_ahead) *schedule_ahead = s->scheduleAheadAmount;
    return ZL_OK;
}

}  // extern "C"
