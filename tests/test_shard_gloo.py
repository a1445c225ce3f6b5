"""Multi-process host logic of the sharded render on CPU: world_size 2, gloo.

Each rank owns the sessions `session mod world == rank`, renders them (here with the CPU oracle standing in for the
device: this is a test of the partition + reduce plumbing, not of the kernels) and the ranks all-reduce their partial
stereo mixes.  The result must equal the serial single-process mix of all sessions within 1e-5 of full scale
(float add reassociation), and every session's own master must be bit-identical to the single-process render."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from libzl_b200 import shard

N_SESS, BLOCKS, FRAMES = 6, 3, 256


def _render_sessions(ids):
    """Render the given global session ids with the oracle; returns master [len(ids), 2, BLOCKS*FRAMES]."""
    from oracle import loader
    from libzl_b200.engine import Engine
    lib = loader.load_oracle()
    e = Engine(lib=lib, n_sessions=len(ids), buses_per_session=2, voices_per_bus=4)
    for local, sid in enumerate(ids):
        rng = np.random.default_rng(100 + sid)                      # everything seeded by the GLOBAL session id
        pcm = rng.uniform(-1, 1, (2, 12000)).astype(np.float32)
        c = e.add_clip(pcm)
        e.set_clip_params(c, attack=0.001, decay=0.01, sustain=0.8, release=0.02, pan=float(rng.uniform(-1, 1)), length_beats=1.5)
        for k in range(3):
            e.play(c, loop=bool(k & 1), midi_channel=k % 2 - 2, note=int(60 + rng.integers(-5, 6)), volume=float(rng.uniform(0.3, 1)), session=local)
    e.render(BLOCKS, FRAMES)
    m = e.master_all().copy()
    e.close()
    return m


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = shard.units_for_rank(N_SESS, rank, world)
        m = _render_sessions(mine)
        partial = torch.from_numpy(m.sum(axis=0, dtype=np.float32))   # this rank's share of the one spanned mix
        shard.reduce_mix(partial, group=None, dst=None)
        if rank == 0:
            np.save(out + ".mix.npy", partial.numpy())
        np.save(f"{out}.r{rank}.npy", m)
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_partition_helpers():
    assert shard.units_for_rank(7, 0, 2) == [0, 2, 4, 6] and shard.units_for_rank(7, 1, 2) == [1, 3, 5]
    seen = sorted(u for r in range(4) for u in shard.units_for_rank(10, r, 4))
    assert seen == list(range(10))
    for u in range(10):
        r, i = shard.owner_of(u, 4)
        assert shard.units_for_rank(10, r, 4)[i] == u
    assert shard.mix_groups(8, 4) == [[0, 1, 2, 3], [4, 5, 6, 7]]
    with pytest.raises(ValueError):
        shard.mix_groups(8, 3)


def test_two_rank_sharded_render_and_mix_reduce(tmp_path):
    world = 2
    out = str(tmp_path / "res")
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    serial = _render_sessions(list(range(N_SESS)))
    # every session renders the same on its owner rank as in the single-process run: sharding is by independent units
    for r in range(world):
        got = np.load(f"{out}.r{r}.npy")
        ids = shard.units_for_rank(N_SESS, r, world)
        assert np.array_equal(got, serial[ids]), f"rank {r}: session masters differ from the single-process render"
    assert np.abs(serial).max() > 0.05, "the scenario must actually sound"
    mix = np.load(out + ".mix.npy")
    ref = np.zeros_like(mix)
    for s in range(N_SESS):                                          # the reference's serial += order over tracks
        ref += serial[s]
    assert np.abs(mix - ref).max() <= 1e-5 * max(1.0, np.abs(ref).max())
