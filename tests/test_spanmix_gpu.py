"""One mix spanning two GPUs through the C ABI (zl_comm_unique_id / zl_comm_init / zl_mix_reduce, NCCL over NVLink):
each rank's own masters bit-exact against the oracle, the reduced mix within 1e-5 of the serial sum (float reassociation),
ncclReduce (root holds the mix, the other rank keeps its partial) and ncclAllReduce.  Needs two devices: skipped otherwise
(run with `gpurun --gpus 2`)."""
import ctypes as C
import multiprocessing as mp
import os
import sys
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
N_SESS, BLOCKS, FRAMES, STEPS = 3, 4, 256, 3


def _session_setup(e, rank):
    for s in range(N_SESS):
        rng = np.random.default_rng(1000 * rank + s)
        c = e.add_clip(rng.uniform(-1, 1, (2, 16000)).astype(np.float32))
        e.set_clip_params(c, attack=0.002, decay=0.01, sustain=0.8, release=0.02, pan=float(rng.uniform(-1, 1)), length_beats=1.5)
        for k in range(4):
            e.play(c, loop=bool(k & 1), midi_channel=k % 2 - 2, note=int(60 + rng.integers(-7, 8)), volume=float(rng.uniform(0.3, 1)), session=s)


def _worker(rank, world, idfile, q):
    try:
        sys.path.insert(0, ROOT)
        import torch  # noqa: F401  (brings the wheel's libnccl.so.2 into the process)
        import libzl_b200
        from libzl_b200.engine import Engine
        lib = libzl_b200.load()
        e = Engine(lib=lib, n_sessions=N_SESS, buses_per_session=2, voices_per_bus=4, device=rank)
        _session_setup(e, rank)
        buf = (C.c_ubyte * 128)()
        if rank == 0:
            e._chk(lib.comm_unique_id(buf), "comm_unique_id")
            with open(idfile + ".tmp", "wb") as f:
                f.write(bytes(buf))
            os.replace(idfile + ".tmp", idfile)
        else:
            for _ in range(600):
                if os.path.exists(idfile):
                    break
                time.sleep(0.05)
            C.memmove(buf, open(idfile, "rb").read(), 128)
        e._chk(lib.comm_init(e.ctx, buf, rank, world), "comm_init")
        own, red = [], []
        for step in range(STEPS):
            e.render(BLOCKS, FRAMES, e.make_clocks(BLOCKS, FRAMES, first_block=step * BLOCKS))
            own.append(e.master_all().copy())
            e._chk(lib.mix_reduce(e.ctx, 0 if step < 2 else -1), "mix_reduce")        # two ncclReduce to rank 0, then one ncclAllReduce
            red.append(e.master_all().copy())
        e._chk(lib.comm_destroy(e.ctx), "comm_destroy")
        e.close()
        q.put((rank, "ok", own, red))
    except Exception:   # noqa: BLE001
        import traceback
        q.put((rank, "error", traceback.format_exc(), None))


def test_span_mix_reduce_over_two_gpus(oracle, tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    from libzl_b200.engine import Engine
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, str(tmp_path / "ncclid"), q)) for r in range(2)]
    for p in procs:
        p.start()
    res = {}
    for _ in range(2):
        rank, status, a, b = q.get(timeout=300)
        assert status == "ok", a
        res[rank] = (a, b)
    for p in procs:
        p.join(timeout=30)
    # per-rank masters: bit-exact against the oracle
    for rank in range(2):
        e = Engine(lib=oracle, n_sessions=N_SESS, buses_per_session=2, voices_per_bus=4)
        _session_setup(e, rank)
        for step in range(STEPS):
            e.render(BLOCKS, FRAMES, e.make_clocks(BLOCKS, FRAMES, first_block=step * BLOCKS))
            assert np.array_equal(e.master_all(), res[rank][0][step]), (rank, step)
        e.close()
    for step in range(STEPS):
        serial = res[0][0][step] + res[1][0][step]
        assert np.abs(serial).max() > 0.05
        assert np.abs(res[0][1][step] - serial).max() <= 1e-5                 # the mix on the root
        if step < 2:
            assert np.array_equal(res[1][1][step], res[1][0][step])           # ncclReduce: the other rank keeps its partial
        else:
            assert np.abs(res[1][1][step] - serial).max() <= 1e-5             # ncclAllReduce: everybody has the mix
