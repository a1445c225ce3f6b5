// shim: lib/JUCEHeaders.h includes this module; nothing on the pinned path uses it
#pragma once
