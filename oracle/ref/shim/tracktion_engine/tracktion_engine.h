// shim: lib/JUCEHeaders.h and lib/Helper.h name these tracktion_engine types in declarations only
#pragma once
#include <juce_core/juce_core.h>
namespace tracktion_engine {
class DeviceManager { public: double getCpuUsage() const { return 0.0; } };
class Engine { public: DeviceManager& getDeviceManager() { static DeviceManager m; return m; } };
class Edit; class AudioTrack; class AudioFile; class AudioFileInfo;
class WaveAudioClip : public juce::ReferenceCountedObject { public: using Ptr = juce::ReferenceCountedObjectPtr<WaveAudioClip>; };
}
