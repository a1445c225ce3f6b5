"""Row n2: the command-scheduling front-end of the product (zl_sched_*, host logic inside
libzl_b200.so - no device needed) against the oracle's restatement, which tests/test_libzl_ref.py pins to the reference's own
compiled lib/SyncTimer.cpp (scenario "scheduler_sketch", live and by digest).  Compared: every period's block clock, every
emitted command (fields, period, tick), the timer state - one period per call and eight periods per call."""
import ctypes as C

import numpy as np
import pytest

from libzl_b200 import _capi as K
import libref as R


def run_scheduler(lib, ops, batch, sample_rate=48000.0, buffer_size=1024):
    s = C.c_void_p()
    assert lib.sched_create(float(sample_rate), buffer_size, C.byref(s)) == K.ZL_OK
    out, period_no = [], 0
    for op in ops:
        if op[0] == "sched_start":
            assert lib.sched_start(s, op[1]) == K.ZL_OK
        elif op[0] == "sched_stop":
            assert lib.sched_stop(s) == K.ZL_OK
        elif op[0] == "sched_bpm":
            assert lib.sched_set_bpm(s, op[1]) == K.ZL_OK
        elif op[0] == "sched_cmd":
            c = K.zl_clip_command()
            lib.clip_command_default(C.byref(c))
            for k, v in op[2].items():
                setattr(c, k, v)
            assert lib.sched_schedule(s, C.byref(c), op[1]) == K.ZL_OK
        elif op[0] == "sched_queue_start":
            assert lib.sched_queue_start(s, op[1], op[2]) == K.ZL_OK
        elif op[0] == "sched_queue_stop":
            assert lib.sched_queue_stop(s, op[1], op[2]) == K.ZL_OK
        elif op[0] == "sched_enable":
            assert lib.sched_schedule_channel_enabled(s, op[1], op[2], op[3]) == K.ZL_OK
        elif op[0] == "sched_run":
            _, nper, nf = op
            period = int(round(nf * 1e6 / sample_rate))
            done = 0
            while done < nper:
                nb = min(batch, nper - done)
                ck = (K.zl_block_clock * nb)()
                for i in range(nb):
                    ck[i].current_usecs = 1_000_000 + (period_no + i) * period
                    ck[i].next_usecs = ck[i].current_usecs + period
                ev = (K.zl_sched_event * 1024)()
                n = C.c_int32()
                assert lib.sched_advance(s, nb, nf, ck, ev, 1024, C.byref(n)) == K.ZL_OK
                for i in range(nb):
                    out.append(("clock", period_no + i, ck[i].playhead, ck[i].playhead_usecs, ck[i].subbeat_usecs))
                for i in range(n.value):
                    x = ev[i]
                    c = x.cmd
                    out.append(("event", period_no + x.block, x.kind, x.tick, x.channel, x.enabled, c.clip, c.midi_note, c.midi_channel, c.start_playback,
                                c.stop_playback, c.change_slice, c.slice, c.change_looping, c.looping, c.change_volume, float(c.volume)))
                period_no += nb
                done += nb
            st = [C.c_uint64() for _ in range(4)] + [C.c_int32(), C.c_uint64()]
            lib.sched_state(s, *[C.byref(x) for x in st])
            out.append(("state",) + tuple(int(x.value) for x in st))
    lib.sched_destroy(s)
    # events are reported after the clocks of their batch: order by period for comparison across batch sizes
    clocks = [x for x in out if x[0] == "clock"]
    events = sorted([x for x in out if x[0] == "event"], key=lambda x: x[1])
    states = [x for x in out if x[0] == "state"]
    return clocks, events, states


@pytest.mark.parametrize("nframes", [64, 256, 512, 1024])
def test_product_scheduler_equals_the_oracle(oracle, nframes):
    import libzl_b200
    product = libzl_b200.load()                      # the scheduler is host code inside the product library: no device needed
    for ops in (R.scenario_scheduler(nframes), R.random_scheduler_ops(nframes, nframes), R.random_scheduler_ops(1000 + nframes, nframes)):
        ref = run_scheduler(oracle, ops, 1)
        assert len(ref[1]) > 3
        for batch in (1, 8):
            got = run_scheduler(product, ops, batch)
            assert got[0] == ref[0], "block clocks differ"
            assert got[1] == ref[1], "emitted commands differ"
            assert got[2] == ref[2], "timer state differs"
