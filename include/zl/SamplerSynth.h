/* SamplerSynth - see libzl_facade.h (mirrors lib/SamplerSynth.h of the reference) */
#include "libzl_facade.h"
