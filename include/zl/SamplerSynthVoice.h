/* SamplerSynthVoice - see libzl_facade.h (mirrors lib/SamplerSynthVoice.h of the reference) */
#include "libzl_facade.h"
