/* ClipAudioSourcePositionsModel - see libzl_facade.h (mirrors lib/ClipAudioSourcePositionsModel.h of the reference) */
#include "libzl_facade.h"
