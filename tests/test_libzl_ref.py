"""Rows a1 / a3 / a5 / a6 / a9 / a10 / a14 pinned: the CPU oracle against the reference's OWN compiled
SamplerSynthVoice::process, SamplerChannel (ring, voice allocation, serial sum), JackPassthrough, positions model and
AudioLevels (built in place into oracle/_ref, see oracle/ref/libzl_ref.cpp) - live where oracle/_ref exists, and through the
committed digests of the reference's outputs (tests/golden/libzl_digests.json) everywhere.  CPU only; the device runs the
same scenarios in tests/test_parity_gpu.py."""
import json
import os

import pytest

import libref as R

GOLDEN = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "libzl_digests.json")))


@pytest.mark.parametrize("name", sorted(R.SCENARIOS))
def test_oracle_matches_reference_digest(oracle, name):
    ops = R.SCENARIOS[name]()
    assert R.digest(R.run_engine(oracle, ops), R.disabled_buses(ops)) == GOLDEN[name], \
        f"{name}: the oracle's output differs from what the reference's own compiled code produced"


@pytest.mark.parametrize("name", sorted(R.SCENARIOS))
def test_oracle_matches_reference_live(oracle, name):
    if not R.available():
        pytest.skip("oracle/_ref/libzl_juceref.so not built (needs /root/reference)")
    ops = R.SCENARIOS[name]()
    ref = R.run_reference(ops)
    assert not R.compare(ref, R.run_engine(oracle, ops), skip_buses=R.disabled_buses(ops))
    assert R.digest(ref, R.disabled_buses(ops)) == GOLDEN[name]
    assert max(float(abs(r["raw"]).max()) for r in ref) > 0.05
