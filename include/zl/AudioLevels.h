/* AudioLevels - see libzl_facade.h (mirrors lib/AudioLevels.h of the reference) */
#include "libzl_facade.h"
