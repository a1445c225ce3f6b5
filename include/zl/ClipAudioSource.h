/* ClipAudioSource - see libzl_facade.h (mirrors lib/ClipAudioSource.h of the reference) */
#include "libzl_facade.h"
