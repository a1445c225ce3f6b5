/* SyncTimer - see libzl_facade.h (mirrors the clip-command half of lib/SyncTimer.h of the reference) */
#include "libzl_facade.h"
