#pragma once
#include "types.h"
#ifdef __cplusplus
extern "C" {
#endif
jack_client_t* jack_client_open(const char* client_name, jack_options_t options, jack_status_t* status, ...);
int jack_client_close(jack_client_t* client);
int jack_set_process_callback(jack_client_t* client, JackProcessCallback cb, void* arg);
int jack_set_xrun_callback(jack_client_t* client, JackXRunCallback cb, void* arg);
int jack_set_buffer_size_callback(jack_client_t* client, JackBufferSizeCallback cb, void* arg);
int jack_set_sample_rate_callback(jack_client_t* client, JackSampleRateCallback cb, void* arg);
int jack_set_port_registration_callback(jack_client_t* client, JackPortRegistrationCallback cb, void* arg);
void jack_on_shutdown(jack_client_t* client, JackShutdownCallback cb, void* arg);
jack_port_t* jack_port_register(jack_client_t* client, const char* port_name, const char* port_type, unsigned long flags, unsigned long buffer_size);
int jack_activate(jack_client_t* client);
int jack_deactivate(jack_client_t* client);
int jack_connect(jack_client_t* client, const char* source_port, const char* destination_port);
int jack_disconnect(jack_client_t* client, const char* source_port, const char* destination_port);
void* jack_port_get_buffer(jack_port_t* port, jack_nframes_t nframes);
const char* jack_port_name(const jack_port_t* port);
int jack_port_flags(const jack_port_t* port);
jack_port_t* jack_port_by_id(jack_client_t* client, jack_port_id_t id);
jack_port_t* jack_port_by_name(jack_client_t* client, const char* name);
const char** jack_get_ports(jack_client_t* client, const char* port_name_pattern, const char* type_name_pattern, unsigned long flags);
void jack_free(void* ptr);
int jack_get_cycle_times(const jack_client_t* client, jack_nframes_t* current_frames, jack_time_t* current_usecs, jack_time_t* next_usecs, float* period_usecs);
float jack_cpu_load(jack_client_t* client);
jack_nframes_t jack_get_sample_rate(jack_client_t* client);
jack_nframes_t jack_get_buffer_size(jack_client_t* client);
jack_time_t jack_get_time(void);
jack_nframes_t jack_frame_time(const jack_client_t* client);
jack_nframes_t jack_last_frame_time(const jack_client_t* client);
jack_time_t jack_frames_to_time(const jack_client_t* client, jack_nframes_t frames);
jack_transport_state_t jack_transport_query(const jack_client_t* client, jack_position_t* pos);
typedef enum JackLatencyCallbackMode { JackCaptureLatency, JackPlaybackLatency } jack_latency_callback_mode_t;
typedef struct _jack_latency_range { jack_nframes_t min, max; } jack_latency_range_t;
typedef void (*JackLatencyCallback)(jack_latency_callback_mode_t mode, void* arg);
int jack_set_latency_callback(jack_client_t* client, JackLatencyCallback cb, void* arg);
void jack_port_get_latency_range(jack_port_t* port, jack_latency_callback_mode_t mode, jack_latency_range_t* range);
#ifdef __cplusplus
}
#endif
