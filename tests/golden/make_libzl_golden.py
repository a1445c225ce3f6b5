"""Generates tests/golden/libzl_digests.json from the REFERENCE'S OWN compiled hot path (oracle/_ref/libzl_juceref.so:
lib/SamplerSynthVoice.cpp, lib/SamplerSynth.cpp, lib/JackPassthrough.cpp, lib/ClipAudioSourcePositionsModel.cpp,
lib/AudioLevels.cpp built in place by `make -C oracle ref`; driver oracle/ref/libzl_ref.cpp, scenarios tests/libref.py).

Run in the build container only (needs /root/reference):  python tests/golden/make_libzl_golden.py
Each scenario's outputs (raw / dry / fx planes of the 12 buses, positions, integer meter peaks) are kept as one sha256, so the
GPU box can check the product against the reference bit for bit without the reference.
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import libref as R  # noqa: E402

if __name__ == "__main__":
    assert R.available(), "oracle/_ref/libzl_juceref.so is missing: make -C oracle ref"
    out = {}
    for name, make in R.SCENARIOS.items():
        ops = make()
        out[name] = R.digest(R.run_reference(ops), R.disabled_buses(ops))
        print(name, out[name])
    with open(os.path.join(HERE, "libzl_digests.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
