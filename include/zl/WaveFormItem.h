/* WaveFormItem - see libzl_facade.h (mirrors lib/WaveFormItem.h of the reference) */
#include "libzl_facade.h"
