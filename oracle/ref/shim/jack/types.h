// oracle/ref/shim/jack — TEST INFRASTRUCTURE: the slice of the JACK C API the reference's hot-path sources call,
// implemented in-process by oracle/ref/libzl_ref.cpp (ports are plain float buffers, the test drives the process cycle).
#pragma once
#include <stdint.h>
#include <stddef.h>
#include <errno.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef uint32_t jack_nframes_t;
typedef uint64_t jack_time_t;
typedef uint64_t jack_uuid_t;
typedef float jack_default_audio_sample_t;
typedef struct _jack_client jack_client_t;
typedef struct _jack_port jack_port_t;
typedef uint32_t jack_port_id_t;
typedef enum JackOptions { JackNullOption = 0, JackNoStartServer = 1 } jack_options_t;
typedef enum JackStatus { JackFailure = 1 } jack_status_t;
enum JackPortFlags { JackPortIsInput = 1, JackPortIsOutput = 2, JackPortIsPhysical = 4, JackPortCanMonitor = 8, JackPortIsTerminal = 16 };
typedef enum { JackTransportStopped = 0, JackTransportRolling = 1, JackTransportLooping = 2, JackTransportStarting = 3 } jack_transport_state_t;
typedef enum { JackPositionBBT = 0x10, JackPositionTimecode = 0x20, JackBBTFrameOffset = 0x40 } jack_position_bits_t;
typedef struct _jack_position {
    uint64_t unique_1; jack_time_t usecs; jack_nframes_t frame_rate; jack_nframes_t frame; jack_position_bits_t valid;
    int32_t bar, beat, tick; double bar_start_tick; float beats_per_bar, beat_type; double ticks_per_beat, beats_per_minute;
    double frame_time, next_time; jack_nframes_t bbt_offset; float audio_frames_per_video_frame; jack_nframes_t video_offset;
    double tick_double; int32_t padding[5]; uint64_t unique_2;
} jack_position_t;
#define JACK_DEFAULT_AUDIO_TYPE "32 bit float mono audio"
#define JACK_DEFAULT_MIDI_TYPE "8 bit raw midi"
typedef int (*JackProcessCallback)(jack_nframes_t nframes, void* arg);
typedef int (*JackXRunCallback)(void* arg);
typedef int (*JackBufferSizeCallback)(jack_nframes_t nframes, void* arg);
typedef int (*JackSampleRateCallback)(jack_nframes_t nframes, void* arg);
typedef void (*JackPortRegistrationCallback)(jack_port_id_t port, int reg, void* arg);
typedef void (*JackShutdownCallback)(void* arg);
#ifdef __cplusplus
}
#endif
