/* JackPassthrough - see libzl_facade.h (mirrors lib/JackPassthrough.h of the reference) */
#include "libzl_facade.h"
