// shim: lib/JUCEHeaders.h includes the whole module (which needs juce_gui_basics); nothing on the pinned path uses it
#pragma once
