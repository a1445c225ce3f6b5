/* ClipCommand - see libzl_facade.h (mirrors lib/ClipCommand.h of the reference) */
#include "libzl_facade.h"
