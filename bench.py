#!/usr/bin/env python
"""bench.py — voice-frames/s of the batched sampler voice render (BASELINE.json metric) on N B200s.

Workload ("cfg2x": BASELINE.json configs[1] shape batched as configs[4]): S independent sessions per GPU,
each a 128-voice polyphonic SamplerSynth (16 buses x 8 voices) over its own 8 stereo 48 kHz 1 s clips
(noise / sweep), random pitch ratios 0.5-2.0, ADSR, M/S pan, clip volume, JackPassthrough, per-bus peak
meters, mixed to one stereo master per session.  One step = one launch of `blocks` JACK periods of 512
frames for every session, preceded by that step's command stream (one-shot re-triggers).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
N > 1 runs under torchrun (one rank per GPU); sessions are sharded by rank with no data-path collective
(weak scaling: sessions per GPU fixed), unless --span-mix adds the NCCL stereo-bus reduce.
Prints ONE JSON line (rank 0).
"""
import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FRAMES = 512
BUSES, VOICES, CLIPS_PER_SESSION, CLIP_FRAMES = 16, 8, 8, 48000
RATE = 48000.0


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--sessions", type=int, default=4096, help="sessions per GPU")
    ap.add_argument("--blocks", type=int, default=16, help="512-frame periods per step")
    ap.add_argument("--cpu-sessions", type=int, default=0, help="sessions of the CPU sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-latency", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle replay of four sessions after the timed regions")
    ap.add_argument("--no-kernel-events", action="store_true", help="experiment: no per-launch event pairs in the timed region")
    ap.add_argument("--rms", action="store_true", help="also produce the per-bus f64 sum of squares (getRMSLevel extension; libzl's AudioLevels has peaks only)")
    ap.add_argument("--no-thumb", action="store_true", help="skip the WaveFormItem thumbnail probe (BASELINE configs[3])")
    ap.add_argument("--span-mix", type=int, default=0, help="G>1: groups of G GPUs share one mix -> NCCL reduce of the masters")
    ap.add_argument("--seed", type=int, default=5)
    ap.add_argument("--quirks", type=int, default=7, help="ZL_Q* bits of the Z-mode voice (7 = faithful to libzl, the default)")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# workload description (shared by the CUDA arm and the CPU arms so both render the same thing)
class Workload:
    def __init__(self, sessions, blocks, steps_total, seed, rank=0):
        self.S, self.blocks, self.steps_total = sessions, blocks, steps_total
        rng = np.random.default_rng(seed * 1000 + rank)
        nclips = sessions * CLIPS_PER_SESSION
        self.clip_seed = seed * 7919 + rank
        self.clip_params = dict(
            attack=rng.uniform(0, 0.05, nclips).astype(np.float32), decay=rng.uniform(0, 0.2, nclips).astype(np.float32),
            sustain=rng.uniform(0.3, 1.0, nclips).astype(np.float32), release=rng.uniform(0.01, 0.2, nclips).astype(np.float32),
            pan=rng.uniform(-1, 1, nclips).astype(np.float32), volume=rng.uniform(0.5, 1.0, nclips).astype(np.float32))
        from libzl_b200.engine import default_commands
        nv = sessions * BUSES * VOICES
        # initial polyphony: 6 looping + 2 one-shot voices per bus
        c = default_commands(nv)
        sess = np.repeat(np.arange(sessions, dtype=np.int32), BUSES * VOICES)
        bus = np.tile(np.repeat(np.arange(BUSES, dtype=np.int32), VOICES), sessions)
        slot = np.tile(np.arange(VOICES, dtype=np.int32), sessions * BUSES)
        c["clip"] = sess * CLIPS_PER_SESSION + rng.integers(0, CLIPS_PER_SESSION, nv)
        c["midi_channel"] = bus - 2
        c["midi_note"] = 60 + rng.integers(-12, 13, nv)
        c["start_playback"] = 1
        c["looping"] = (slot < 6).astype(np.int32)
        c["change_volume"] = 1
        c["volume"] = rng.uniform(0.25, 1.0, nv).astype(np.float32)
        # distinct notes per bus slot so commands stay non-equivalent (ClipCommand::equivalentTo)
        self.init = (sess, np.zeros(nv, np.int32), np.zeros(nv, np.uint64), c)
        # per-step command stream: one one-shot start per bus at a random period of the step; it takes the
        # first free voice of the bus or is dropped when all 8 are busy (lib/SamplerSynth.cpp:204-215)
        self.steps = []
        nb = sessions * BUSES
        for _ in range(steps_total):
            k = default_commands(nb)
            s2 = np.repeat(np.arange(sessions, dtype=np.int32), BUSES)
            k["clip"] = s2 * CLIPS_PER_SESSION + rng.integers(0, CLIPS_PER_SESSION, nb)
            k["midi_channel"] = np.tile(np.arange(BUSES, dtype=np.int32), sessions) - 2
            k["midi_note"] = 60 + rng.integers(-12, 13, nb)
            k["start_playback"] = 1
            k["change_volume"] = 1
            k["volume"] = rng.uniform(0.25, 1.0, nb).astype(np.float32)
            self.steps.append((s2, rng.integers(0, blocks, nb).astype(np.int32), np.zeros(nb, np.uint64), k))

    def clip_pcm_numpy(self, first, count):
        """Interleaved [count, CLIP_FRAMES + 8, 2] float32 (zero padded) for clips first..first+count-1."""
        out = np.zeros((count, CLIP_FRAMES + 8, 2), np.float32)
        t = np.arange(CLIP_FRAMES) / RATE
        ph = 2 * np.pi * (20.0 * t + (20000.0 - 20.0) * t * t / 2.0)
        for i in range(count):
            cid = first + i
            if cid % 2 == 0:
                out[i, :CLIP_FRAMES] = np.random.default_rng(self.clip_seed * 100003 + cid).uniform(-1, 1, (CLIP_FRAMES, 2))
            else:
                g = 0.4 + 0.4 * ((cid * 2654435761) % 1000) / 1000.0
                out[i, :CLIP_FRAMES, 0] = g * np.sin(ph)
                out[i, :CLIP_FRAMES, 1] = g * np.sin(ph + np.pi / 2)
        return out

    def configure_clip(self, e, cid, as_id=None):
        p = self.clip_params
        e.set_clip_params(cid if as_id is None else as_id, attack=float(p["attack"][cid]), decay=float(p["decay"][cid]),
                          sustain=float(p["sustain"][cid]), release=float(p["release"][cid]), pan=float(p["pan"][cid]),
                          volume_abs=float(p["volume"][cid]), length_beats=1.5, root_note=60)

    def subset(self, batch, sessions):
        """The commands of `batch` = (sessions, blocks, ticks, cmds) that belong to the given global sessions, renumbered to
        local sessions 0.. and local clip ids (a session's 8 clips are private to it)."""
        sess, blocks, ticks, c = batch
        local = {int(g): i for i, g in enumerate(sessions)}
        keep = np.isin(sess, list(local))
        s2 = np.array([local[int(x)] for x in sess[keep]], np.int32)
        c2 = c[keep].copy()
        c2["clip"] = s2 * CLIPS_PER_SESSION + (c2["clip"] - sess[keep] * CLIPS_PER_SESSION)
        return s2, blocks[keep], ticks[keep], c2


def clocks_for(e, step, blocks):
    return e.make_clocks(blocks, FRAMES, RATE, first_block=step * blocks)


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe: clocks.sm, clocks.max.sm and
    the clocks_event_reasons of nvidia-smi).  Read through NVML in this process every 5 ms - `nvidia-smi -lms` needs a few
    hundred milliseconds to print its first line, longer than a ten-step timed region - with the nvidia-smi loop as fallback."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    BITS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def __init__(self, index):
        self.rows, self.proc, self.nvml, self.stopFlag, self.source = [], None, None, False, None
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            try:
                index = int(vis.split(",")[index])
            except Exception:
                pass
        try:
            import pynvml
            pynvml.nvmlInit()
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.nvml, self.source = pynvml, "nvml, every 5 ms"
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.source = "nvidia-smi -lms 100"
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _poll(self):
        n = self.nvml
        reasons = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self.stopFlag:
            try:
                sm = float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM))
                mask = int(reasons(self.h))
                self.rows.append([sm, self.mx, 0.0] + ["Active" if mask & bit else "Not Active" for _, bit in self.BITS])
            except Exception:
                pass
            time.sleep(0.005)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.nvml:
            self.stopFlag = True
            self.t.join(timeout=1)
        elif self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        else:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
                for (n, _), v in zip(self.BITS, r[3:7]):
                    if str(v).lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        busy = [x for x in sm if x > 0]
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": self.source}


def S_BUS_BYTES(args):
    return args.sessions * BUSES * 2 * args.blocks * FRAMES * 4      # raw bus planes of one step on one GPU


TRAFFIC_FILE = os.path.join("profiles", "r02_traffic.json")


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of one gather_kernel launch of this workload.  NOT measured by this run:
    read from the committed summary of a builder-run `ncu --set full` capture of the same command (roofline.traffic_source
    says so); None when the file is missing or the workload differs."""
    try:
        with open(os.path.join(ROOT, TRAFFIC_FILE)) as f:
            t = json.load(f)["gather"]
        return t["dram_bytes_read"] + t["dram_bytes_write"]
    except Exception:
        return None


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, torch copy read+write)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def run_cuda(args, rank, world, local_rank):
    import torch
    import libzl_b200
    from libzl_b200 import _capi as K
    from libzl_b200.engine import Engine
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    S, NB = args.sessions, args.blocks
    total_steps = args.warmup + args.steps
    wl = Workload(S, NB, total_steps * 2 + 4 + args.steps, args.seed, rank)
    outputs = K.ZL_OUT_RAW | K.ZL_OUT_MASTER | K.ZL_OUT_LEVELS | (K.ZL_OUT_RMS if args.rms else 0)
    e = Engine(lib=libzl_b200.load(), n_sessions=S, buses_per_session=BUSES, voices_per_bus=VOICES,
               max_clips=S * CLIPS_PER_SESSION, max_frames=FRAMES, max_blocks=NB, outputs=outputs, device=local_rank,
               quirks=args.quirks)
    stream = torch.cuda.Stream(device=dev)          # a real (non-default) stream: the context launches on exactly this one
    e.set_stream(stream.cuda_stream)
    # clips: generated on the device (synthetic noise / sweep), adopted zero-copy by the clip table
    nclips = S * CLIPS_PER_SESSION
    g = torch.Generator(device=dev); g.manual_seed(wl.clip_seed)
    pcm = torch.zeros((nclips, CLIP_FRAMES + 8, 2), dtype=torch.float32, device=dev)
    chunk = 1024
    t = torch.arange(CLIP_FRAMES, device=dev, dtype=torch.float64) / RATE
    ph = 2 * torch.pi * (20.0 * t + (20000.0 - 20.0) * t * t / 2.0)
    sweep = torch.stack([torch.sin(ph), torch.sin(ph + torch.pi / 2)], dim=1).to(torch.float32)
    for c0 in range(0, nclips, chunk):
        n = min(chunk, nclips - c0)
        blk = pcm[c0:c0 + n, :CLIP_FRAMES]
        blk.uniform_(-1.0, 1.0, generator=g)
        gains = (0.4 + 0.4 * torch.rand(n, generator=g, device=dev)).view(n, 1, 1)
        odd = (torch.arange(c0, c0 + n, device=dev) % 2 == 1).view(n, 1, 1)
        blk.copy_(torch.where(odd, gains * sweep.unsqueeze(0), blk))
    torch.cuda.synchronize()
    base = pcm.data_ptr()
    clip_bytes = (CLIP_FRAMES + 8) * 2 * 4
    for cid in range(nclips):
        got = e.add_clip_device(base + cid * clip_bytes, 2, CLIP_FRAMES, RATE)
        assert got == cid
        wl.configure_clip(e, cid)
    e.command_batch(*wl.init)

    n_per_step = NB * FRAMES
    out_bytes = (S * BUSES * 2 * n_per_step * 4) + (S * 2 * n_per_step * 4) + NB * S * BUSES * 2 * 4
    state_bytes = 2 * S * BUSES * VOICES * 224

    # --span-mix G: groups of G ranks share one stereo mix per session index -> NCCL sum of the master planes, in place on
    # the library's own device buffer, inside the timed region (the only collective on the path)
    # --span-mix G: groups of G ranks share one stereo mix per session index -> the library's own collective
    # (zl_comm_init / zl_mix_reduce: ncclReduce of the master planes to the group's first rank, in place, on the context
    # stream), inside the timed region (the only collective on the path).  Only the group's root reads the mix back.
    span, mix_root = False, True
    if args.span_mix > 1 and dist is not None:
        from libzl_b200 import shard
        for ranks in shard.mix_groups(world, args.span_mix):
            gobj = dist.new_group(ranks=ranks)
            if rank in ranks:
                idt = torch.zeros(128, dtype=torch.uint8, device=dev)
                if rank == ranks[0]:
                    buf = (C.c_ubyte * 128)()
                    e._chk(e.lib.comm_unique_id(buf), "comm_unique_id")
                    idt.copy_(torch.tensor(list(bytes(buf)), dtype=torch.uint8))
                dist.broadcast(idt, src=ranks[0], group=gobj)
                buf = (C.c_ubyte * 128)(*idt.cpu().tolist())
                e._chk(e.lib.comm_init(e.ctx, buf, ranks.index(rank), len(ranks)), "comm_init")
                span, mix_root = True, rank == ranks[0]

    def step(i, read_back=None):
        e.command_batch(*wl.steps[i])
        e.render(NB, FRAMES, clocks_for(e, i, NB))
        if span:
            e._chk(e.lib.mix_reduce(e.ctx, 0), "mix_reduce")
        if read_back is not None:
            e._chk(e.lib.read_master_all(e.ctx, read_back, n_per_step), "read_master_all")

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing ("value") -----------------------------------------------------
    for i in range(args.warmup):
        step(i)
    barrier()
    # gather_kernel alone (no planner beside it): one extra untimed step on one stream, per-launch events
    e._chk(e.lib.ctx_set_overlap(e.ctx, 0), "ctx_set_overlap")
    step(2 * total_steps)                                     # one step to settle into the single-stream schedule
    barrier()
    e._chk(e.lib.kernel_timing(e.ctx, 2), "kernel_timing")
    step(2 * total_steps + 1)
    barrier()
    kms0, kn0 = C.c_double(), C.c_uint64()
    e._chk(e.lib.kernel_time(e.ctx, C.byref(kms0), C.byref(kn0)), "kernel_time")
    e._chk(e.lib.kernel_timing(e.ctx, 0), "kernel_timing")
    e._chk(e.lib.ctx_set_overlap(e.ctx, 1), "ctx_set_overlap")
    gather_alone_ms = kms0.value / max(1, kn0.value)
    sampler = ClockSampler(local_rank) if rank == 0 and not os.environ.get("ZL_BENCH_NO_SAMPLER") else None
    launches0 = e.launch_count()
    vf0, src0 = e.total_counters()
    e._chk(e.lib.kernel_timing(e.ctx, 0 if args.no_kernel_events else 2), "kernel_timing")    # CUDA events around each gather_kernel launch, same stream
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    e.lib.host_profile(e.ctx, None, None, None)                         # clear the host-time sums
    ev[0].record(stream)
    th0 = time.perf_counter()
    for k in range(args.steps):
        step(args.warmup + k)
        ev[k + 1].record(stream)
    host_loop_ms = (time.perf_counter() - th0) * 1e3 / args.steps        # wall time of the submitting loop (includes back-pressure)
    barrier()
    hs, hw, hn = C.c_double(), C.c_double(), C.c_uint64()
    e._chk(e.lib.host_profile(e.ctx, C.byref(hs), C.byref(hw), C.byref(hn)), "host_profile")
    host_submit_ms = (hs.value - hw.value) / max(1, hn.value)           # host work inside zl_render per step, without the wait for a free command slot
    elapsed_ms = ev[0].elapsed_time(ev[-1])
    kernel_launches = e.launch_count() - launches0
    kms, kn = C.c_double(), C.c_uint64()
    e._chk(e.lib.kernel_time(e.ctx, C.byref(kms), C.byref(kn)), "kernel_time")
    e._chk(e.lib.kernel_timing(e.ctx, 0), "kernel_timing")
    kernel_ms = kms.value / max(1, kn.value)                  # mean duration of one gather_kernel launch (one period)
    gather_ms_per_step = kms.value / args.steps
    clocks = sampler.stop() if sampler else None
    step_ms = [ev[k].elapsed_time(ev[k + 1]) for k in range(args.steps)]
    vf1, src1 = e.total_counters()
    timed_vf, timed_src = vf1 - vf0, src1 - src0
    # ---- e2e: same steps through the C ABI with host buffers (commands from pinned host memory H2D, master D2H)
    # two pinned host buffers: the D2H copy of step k's masters overlaps the render of step k + 1 (zl_read_master_all_async);
    # every byte of every step's result is on the host when the timed region ends (zl_read_wait before the closing barrier)
    hostp = [C.c_void_p(), C.c_void_p()]
    for hp in hostp:
        e._chk(e.lib.host_alloc(C.byref(hp), S * 2 * n_per_step * 4), "host_alloc")
    host_master = [C.cast(hp, C.POINTER(C.c_float)) for hp in hostp]
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        i = total_steps + k
        e.command_batch(*wl.steps[i])
        e.render(NB, FRAMES, clocks_for(e, i, NB))
        if span:
            e._chk(e.lib.mix_reduce(e.ctx, 0), "mix_reduce")
        if mix_root:                                           # a spanned mix lives on its group's root only
            e._chk(e.lib.read_master_all_async(e.ctx, host_master[k & 1], n_per_step), "read_master_all_async")
    e._chk(e.lib.read_wait(e.ctx), "read_wait")
    barrier()
    e2e_s = time.perf_counter() - t0
    e2e_vf = e.total_counters()[0] - vf1
    h2d = wl.steps[0][3].nbytes + wl.steps[0][0].nbytes * 2 + wl.steps[0][2].nbytes + NB * 40
    d2h = S * 2 * n_per_step * 4 if mix_root else 0
    # the last step's masters really are on the host: compare a few sessions with a synchronous read-back
    if mix_root:
        chk = np.ctypeslib.as_array(host_master[(args.steps - 1) & 1], shape=(S, 2, n_per_step))[:4].copy()
        for sidx in range(min(4, S)):
            l_, r_ = e.master(sidx)
            assert np.array_equal(chk[sidx, 0], l_) and np.array_equal(chk[sidx, 1], r_), "pipelined read-back differs from the synchronous one"
    # ---- parity at the benchmarked size: four of the S sessions replayed through the CPU oracle (outside every timed region)
    parity = None
    if not span and not args.no_parity:
        try:
            order = list(range(args.warmup)) + [2 * total_steps, 2 * total_steps + 1] + list(range(args.warmup, total_steps)) \
                + list(range(total_steps, total_steps + args.steps))
            sel = sorted({0, S // 3, (2 * S) // 3, S - 1})
            last = np.ctypeslib.as_array(host_master[(args.steps - 1) & 1], shape=(S, 2, n_per_step))[sel].copy()
            pk_all, _ = e.block_levels()
            vinfo = e.voices()
            clips_host = [pcm[s_ * CLIPS_PER_SESSION:(s_ + 1) * CLIPS_PER_SESSION, :CLIP_FRAMES].cpu().numpy() for s_ in sel]
            parity = parity_at_size(args, wl, sel, order, clips_host, last, pk_all, vinfo)
        except Exception as ex:   # noqa: BLE001
            parity = {"error": repr(ex)}
    for hp in hostp:
        e.lib.host_free(hp)
    # ---- e2e with the masters read back as 16-bit PCM (the reference recorder's file format, encoded on the device by the
    # same conversion zl_bounce_wav uses: half the D2H bytes).  Same loop, same rules; the f32 loop above stays the parity path.
    pcm_bytes = S * n_per_step * 2 * 2
    pcmp = [C.c_void_p(), C.c_void_p()]
    for hp in pcmp:
        e._chk(e.lib.host_alloc(C.byref(hp), pcm_bytes), "host_alloc")
    for k in range(2):                                         # untimed: both device-side PCM buffers get allocated
        i = 2 * total_steps + 2 + k
        e.command_batch(*wl.steps[i])
        e.render(NB, FRAMES, clocks_for(e, i, NB))
        if span:
            e._chk(e.lib.mix_reduce(e.ctx, 0), "mix_reduce")
        if mix_root:
            e._chk(e.lib.read_master_all_pcm_async(e.ctx, pcmp[k & 1], n_per_step, 16), "read_master_all_pcm_async")
    e._chk(e.lib.read_wait(e.ctx), "read_wait")
    barrier()
    vf2 = e.total_counters()[0]
    t0 = time.perf_counter()
    for k in range(args.steps):
        i = 2 * total_steps + 4 + k
        e.command_batch(*wl.steps[i])
        e.render(NB, FRAMES, clocks_for(e, i, NB))
        if span:
            e._chk(e.lib.mix_reduce(e.ctx, 0), "mix_reduce")
        if mix_root:
            e._chk(e.lib.read_master_all_pcm_async(e.ctx, pcmp[k & 1], n_per_step, 16), "read_master_all_pcm_async")
    e._chk(e.lib.read_wait(e.ctx), "read_wait")
    barrier()
    e2e_pcm_s = time.perf_counter() - t0
    e2e_pcm_vf = e.total_counters()[0] - vf2
    if mix_root:                                               # the bytes on the host are the data chunk of zl_bounce_wav's file, session by session
        got = np.ctypeslib.as_array(C.cast(pcmp[(args.steps - 1) & 1], C.POINTER(C.c_uint8)), shape=(S, n_per_step * 4))
        for sidx in sorted({0, S // 2, S - 1}):
            wav = e.bounce_wav(sidx, -1, 0, 16)
            assert np.array_equal(got[sidx], np.frombuffer(wav, np.uint8)[104:104 + n_per_step * 4]), "PCM read-back differs from zl_bounce_wav"
    for hp in pcmp:
        e.lib.host_free(hp)
    return dict(e=e, parity=parity, e2e_pcm_s=e2e_pcm_s, e2e_pcm_vf=e2e_pcm_vf, d2h_pcm=pcm_bytes if mix_root else 0, host_loop_ms=host_loop_ms, host_submit_ms=host_submit_ms, kernel_ms=kernel_ms, gather_ms_per_step=gather_ms_per_step, gather_alone_ms=gather_alone_ms, elapsed_ms=elapsed_ms, step_ms=step_ms, timed_vf=timed_vf, timed_src=timed_src, launches=kernel_launches,
                clocks=clocks, e2e_s=e2e_s, e2e_vf=e2e_vf, h2d=h2d, d2h=d2h, out_bytes=out_bytes, state_bytes=state_bytes,
                dist=dist, dev=dev, wl=wl, keep=pcm)


# ------------------------------------------------------------------------------------------------
def parity_at_size(args, wl, sel, order, clips_host, last_master, peaks_all, voices):
    """The sessions `sel` of the benchmarked workload, replayed through the exact CPU oracle (oracle/liboracle.so,
    -ffp-contract=off) over every step the device ran, in the same order; the LAST step's masters (as they arrived on the
    host through the pipelined read-back), its per-period integer peaks and the voice table must be bit-equal."""
    from oracle import loader
    from libzl_b200 import _capi as K
    from libzl_b200.engine import Engine
    t0 = time.perf_counter()
    NB = args.blocks
    o = Engine(lib=loader.load_oracle(), n_sessions=len(sel), buses_per_session=BUSES, voices_per_bus=VOICES,
               max_clips=len(sel) * CLIPS_PER_SESSION, max_frames=FRAMES, max_blocks=NB,
               outputs=K.ZL_OUT_RAW | K.ZL_OUT_MASTER | K.ZL_OUT_LEVELS, quirks=args.quirks)
    for i, s_ in enumerate(sel):
        for k in range(CLIPS_PER_SESSION):
            cid = o.add_clip(np.ascontiguousarray(clips_host[i][k].T), RATE)
            wl.configure_clip(o, s_ * CLIPS_PER_SESSION + k, as_id=cid)
    o.command_batch(*wl.subset(wl.init, sel))
    vf = 0
    for i in order:
        o.command_batch(*wl.subset(wl.steps[i], sel))
        o.render(NB, FRAMES, clocks_for(o, i, NB))
        vf += o.voice_frames()
    m = o.master_all()
    pk, _ = o.block_levels()
    ov = o.voices()
    o.close()
    masters_ok = bool(np.array_equal(m, last_master))
    peaks_ok = all(np.array_equal(pk[:, i * BUSES:(i + 1) * BUSES], peaks_all[:, s_ * BUSES:(s_ + 1) * BUSES]) for i, s_ in enumerate(sel))
    nv = BUSES * VOICES
    voices_ok = True
    for i, s_ in enumerate(sel):
        for k in range(nv):
            a, b = ov[i * nv + k], voices[s_ * nv + k]
            if (a.position, a.playing, a.envelope, a.adsr_state, a.pos_gain) != (b.position, b.playing, b.envelope, b.adsr_state, b.pos_gain):
                voices_ok = False
    res = {"sessions_checked": [int(x) for x in sel], "steps_replayed": len(order), "oracle_voice_frames": int(vf),
           "masters_bit_equal": masters_ok, "peaks_bit_equal": bool(peaks_ok), "voice_table_bit_equal": bool(voices_ok),
           "max_abs_master_diff": float(np.abs(m - last_master).max()), "peak_master": float(np.abs(m).max()),
           "oracle_s": round(time.perf_counter() - t0, 2),
           "what": "exact CPU oracle (-ffp-contract=off) over every step of this run for these sessions, vs the last step's masters on the host, its per-period peaks and the device voice table"}
    assert masters_ok and peaks_ok and voices_ok, f"parity at size FAILED: {res}"
    return res


def run_cpu(args, sessions, steps, warmup, threads=None):
    """The reference's CPU implementation of the path (oracle port of lib/SamplerSynthVoice.cpp::process etc.,
    built like the reference with contraction on: oracle/liboracle_fma.so) on a bounded sample of the workload."""
    from oracle import loader
    from libzl_b200 import _capi as K
    from libzl_b200.engine import Engine
    lib = loader.load_oracle(fma=True)
    threads = threads or os.cpu_count() or 1
    NB = args.blocks
    wl = Workload(sessions, NB, warmup + steps + 1, args.seed, rank=0)
    outputs = K.ZL_OUT_RAW | K.ZL_OUT_MASTER | K.ZL_OUT_LEVELS | (K.ZL_OUT_RMS if args.rms else 0)
    e = Engine(lib=lib, n_sessions=sessions, buses_per_session=BUSES, voices_per_bus=VOICES,
               max_clips=sessions * CLIPS_PER_SESSION, max_frames=FRAMES, max_blocks=NB, outputs=outputs, quirks=args.quirks)
    lib.dll.zlo_set_threads(e.ctx, threads)
    nclips = sessions * CLIPS_PER_SESSION
    for c0 in range(0, nclips, 64):
        n = min(64, nclips - c0)
        pcm = wl.clip_pcm_numpy(c0, n)
        for i in range(n):
            cid = e.add_clip(np.ascontiguousarray(pcm[i, :CLIP_FRAMES].T), RATE)
            wl.configure_clip(e, cid)
    e.command_batch(*wl.init)
    for i in range(warmup):
        e.command_batch(*wl.steps[i]); e.render(NB, FRAMES, clocks_for(e, i, NB))
    vf0 = e.total_counters()[0]
    t0 = time.perf_counter()
    for k in range(steps):
        e.command_batch(*wl.steps[warmup + k]); e.render(NB, FRAMES, clocks_for(e, warmup + k, NB))
    dt = time.perf_counter() - t0
    vf = e.total_counters()[0] - vf0
    e.close()
    return vf, dt, threads


def cpu_sample_size(args):
    if args.cpu_sessions:
        return args.cpu_sessions
    return max(16, 16 * (os.cpu_count() or 1))


def latency_probe(local_rank, iters=1500):
    """p50 of one 512-frame period of ONE 128-voice session (configs[1]): host call -> master on the host."""
    import libzl_b200
    from libzl_b200 import _capi as K
    from libzl_b200.engine import Engine
    wl = Workload(1, 1, 2 * (iters + 20), 99, 0)
    e = Engine(lib=libzl_b200.load(), n_sessions=1, buses_per_session=BUSES, voices_per_bus=VOICES, max_clips=8,
               max_frames=FRAMES, max_blocks=1, outputs=K.ZL_OUT_MASTER | K.ZL_OUT_LEVELS, device=local_rank)
    pcm = wl.clip_pcm_numpy(0, CLIPS_PER_SESSION)
    for i in range(CLIPS_PER_SESSION):
        cid = e.add_clip(np.ascontiguousarray(pcm[i, :CLIP_FRAMES].T), RATE)
        wl.configure_clip(e, cid)
    e.command_batch(*wl.init)
    hostp = C.c_void_p()
    e._chk(e.lib.host_alloc(C.byref(hostp), 2 * FRAMES * 4), "host_alloc")
    hp = C.cast(hostp, C.POINTER(C.c_float))
    out = {}
    step = 0
    for name, graph in (("graph", 1), ("plain_launches", 0)):           # SURVEY.md 8d defines the metric as graph-launched; the other for comparison
        e._chk(e.lib.ctx_set_graph(e.ctx, graph), "ctx_set_graph")
        ts = []
        for i in range(iters + 20):
            clk = clocks_for(e, step, 1)
            step += 1
            t0 = time.perf_counter()
            e.lib.render(e.ctx, 1, FRAMES, clk)
            e.lib.read_master_all(e.ctx, hp, FRAMES)
            ts.append(time.perf_counter() - t0)
        ts = sorted(ts[20:])
        out[name] = {"p50_us": round(1e6 * ts[len(ts) // 2], 2), "p99_us": round(1e6 * ts[int(len(ts) * 0.99)], 2)}
    e.lib.host_free(hostp)
    e.close()
    return {"p50_us": out["graph"]["p50_us"], "p99_us": out["graph"]["p99_us"], "plain_launches": out["plain_launches"],
            "what": "one 512-frame period of one 128-voice session: zl_render (one CUDA-graph launch: counter resets, plan, walk, gather, port mix) "
                    "+ zl_read_master_all to pinned host; plain_launches = the same with the graph switched off (zl_ctx_set_graph 0)"}


def sched_probe(steps=200):
    """Row n2: the command scheduler (host code inside the product library) feeding one session - 96 commands scheduled per
    step onto a running timer, zl_sched_advance over the step's 16 periods."""
    import libzl_b200
    from libzl_b200 import _capi as K
    lib = libzl_b200.load()
    s = C.c_void_p()
    assert lib.sched_create(RATE, FRAMES, C.byref(s)) == K.ZL_OK
    lib.sched_start(s, 120)
    c = K.zl_clip_command()
    lib.clip_command_default(C.byref(c))
    rng = np.random.default_rng(1)
    nb = 16
    period = int(round(FRAMES * 1e6 / RATE))
    ck = (K.zl_block_clock * nb)()
    ev = (K.zl_sched_event * 8192)()
    n = C.c_int32()
    t_sched = t_adv = 0.0
    ncmd = nev = per = 0
    for _ in range(steps):
        t0 = time.perf_counter()
        for k in range(96):
            c.clip, c.midi_channel, c.midi_note, c.start_playback, c.change_volume, c.volume = k % 8, k % 12 - 2, 48 + k // 12, 1, 1, 0.5
            lib.sched_schedule(s, C.byref(c), int(rng.integers(0, 64)))
        t1 = time.perf_counter()
        for i in range(nb):
            ck[i].current_usecs = 1_000_000 + (per + i) * period
            ck[i].next_usecs = ck[i].current_usecs + period
        lib.sched_advance(s, nb, FRAMES, ck, ev, 8192, C.byref(n))
        t2 = time.perf_counter()
        per += nb; ncmd += 96; nev += n.value; t_sched += t1 - t0; t_adv += t2 - t1
    lib.sched_destroy(s)
    return {"advance_us_per_step": round(t_adv / steps * 1e6, 1), "commands_scheduled": ncmd, "commands_emitted": nev,
            "schedule_calls_per_s_through_ctypes": round(ncmd / t_sched),
            "what": "zl_sched_*: 96 commands scheduled per step onto a running 120 bpm timer, zl_sched_advance over 16 periods x 512 frames; one host core"}


def cpu_latency_probe(iters=300):
    """The same single period of one 128-voice session on ONE host thread with the oracle port (contraction on, as the
    reference is built): the CPU figure that stands beside the device latency."""
    from oracle import loader
    from libzl_b200 import _capi as K
    from libzl_b200.engine import Engine
    wl = Workload(1, 1, iters + 20, 99, 0)
    lib = loader.load_oracle(fma=True)
    e = Engine(lib=lib, n_sessions=1, buses_per_session=BUSES, voices_per_bus=VOICES, max_clips=8, max_frames=FRAMES, max_blocks=1,
               outputs=K.ZL_OUT_MASTER | K.ZL_OUT_LEVELS)
    lib.dll.zlo_set_threads(e.ctx, 1)
    pcm = wl.clip_pcm_numpy(0, CLIPS_PER_SESSION)
    for i in range(CLIPS_PER_SESSION):
        cid = e.add_clip(np.ascontiguousarray(pcm[i, :CLIP_FRAMES].T), RATE)
        wl.configure_clip(e, cid)
    e.command_batch(*wl.init)
    out = np.zeros((1, 2, FRAMES), np.float32)
    outp = out.ctypes.data_as(C.POINTER(C.c_float))
    ts = []
    for i in range(iters + 20):
        clk = clocks_for(e, i, 1)
        t0 = time.perf_counter()
        e.lib.render(e.ctx, 1, FRAMES, clk)
        e.lib.read_master_all(e.ctx, outp, FRAMES)
        ts.append(time.perf_counter() - t0)
    e.close()
    ts = sorted(ts[20:])
    return {"p50_us": round(1e6 * ts[len(ts) // 2], 2), "p99_us": round(1e6 * ts[int(len(ts) * 0.99)], 2), "threads": 1,
            "what": "the same period on one host thread: oracle port (-O3 -mfma), render + master copy"}


def thumbnail_probe(local_rank, peak):
    """BASELINE configs[3]: WaveFormItem / juce::AudioThumbnail min/max over a 1 h stereo 48 kHz clip at 4096 pixels.
    level 0 (512-sample int8 min/max blocks) is the HBM-bound kernel: algorithmic bytes = nch * 4 * frames read."""
    import torch
    import libzl_b200
    from libzl_b200 import _capi as K
    from libzl_b200.engine import Engine
    dev = torch.device("cuda", local_rank)
    frames = 3600 * 48000
    e = Engine(lib=libzl_b200.load(), n_sessions=1, buses_per_session=1, voices_per_bus=1, max_clips=2, outputs=K.ZL_OUT_RAW, device=local_rank)
    stream = torch.cuda.Stream(device=dev)
    e.set_stream(stream.cuda_stream)
    g = torch.Generator(device=dev); g.manual_seed(4)
    pcm = torch.zeros((frames + 8, 2), dtype=torch.float32, device=dev)
    pcm[:frames].uniform_(-1.0, 1.0, generator=g)
    t = torch.arange(frames, device=dev, dtype=torch.float32) / 48000.0
    pcm[:frames] *= (0.55 + 0.45 * torch.sin(2 * torch.pi * 0.1 * t)).unsqueeze(1)          # slow 0.1 Hz amplitude envelope
    pcm[::1000003] = 1.0; pcm[500001::1000003] = -1.0                                        # sparse full-scale clicks
    pcm[frames:] = 0.0
    del t
    torch.cuda.synchronize()
    cid = e.add_clip_device(pcm.data_ptr(), 2, frames, 48000.0)
    ms = []
    for _ in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        e.thumb_build(cid, 512)
        e1.record(stream)
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    t0 = time.perf_counter()
    px = e.thumb_pixels(cid, 0.0, 3600.0, 4096)
    px_ms = (time.perf_counter() - t0) * 1e3
    t0 = time.perf_counter()
    mask = e.thumb_raster(cid, 0, 0.0, 3600.0, 4096, 400)
    raster_ms = (time.perf_counter() - t0) * 1e3
    best = min(ms[1:])
    alg = 2 * 4 * frames
    e.close()
    return {"what": "juce::AudioThumbnail level 0 (512-sample int8 min/max) over a 1 h stereo 48 kHz clip, HBM-resident; then 4096 pixels and the 4096x400 column mask to the host",
            "level0_ms": round(best, 4), "algorithmic_bytes": alg, "achieved_GBps": alg / (best * 1e-3) / 1e9, "frac": alg / (best * 1e-3) / 1e9 / peak,
            "pixels_4096_host_ms": round(px_ms, 3), "raster_4096x400_host_ms": round(raster_ms, 3),
            "pixel_checksum": int(px.astype(np.int64).sum()), "mask_on": int(mask.sum()) if mask is not None else None}


def wav_io_probe(local_rank):
    """Rows n1 / n3: a 16-bit stereo WAV image in pinned host memory -> HBM clip (H2D of the file bytes + device decode),
    and host floats -> 16-bit WAV image (H2D + device convert + D2H).  End-to-end times through the C ABI; the kernels'
    own durations are in profiles/ (ncu launch list)."""
    import libzl_b200
    from libzl_b200 import _capi as K
    from libzl_b200.engine import Engine
    frames = 1 << 25                                          # 128 MB of file, 256 MB of clip
    e = Engine(lib=libzl_b200.load(), n_sessions=1, buses_per_session=1, voices_per_bus=1, max_clips=8, outputs=K.ZL_OUT_RAW, device=local_rank)
    nbytes = 44 + frames * 4
    hp = C.c_void_p()
    e._chk(e.lib.host_alloc(C.byref(hp), nbytes), "host_alloc")
    img = np.ctypeslib.as_array(C.cast(hp, C.POINTER(C.c_uint8)), shape=(nbytes,))
    import struct
    img[:44] = np.frombuffer(b"RIFF" + struct.pack("<I", nbytes - 8) + b"WAVEfmt " + struct.pack("<IHHIIHH", 16, 1, 2, 48000, 192000, 4, 16)
                             + b"data" + struct.pack("<I", frames * 4), np.uint8)
    rng = np.random.default_rng(11)
    img[44:] = rng.integers(0, 256, frames * 4, dtype=np.uint8)
    ms = []
    for _ in range(4):
        cid = C.c_int32(-1)
        t0 = time.perf_counter()
        e._chk(e.lib.clip_create_from_wav(e.ctx, hp, nbytes, C.byref(cid)), "clip_create_from_wav")
        ms.append((time.perf_counter() - t0) * 1e3)
        e._chk(e.lib.clip_destroy(e.ctx, cid.value), "clip_destroy")
    ingest = min(ms[1:])
    n = 1 << 24
    x = rng.uniform(-1, 1, (2, n)).astype(np.float32)
    t_enc = []
    for _ in range(3):
        t0 = time.perf_counter()
        out = e.wav_encode(x[0], x[1], rate=48000.0, bits=16)
        t_enc.append((time.perf_counter() - t0) * 1e3)
    e.lib.host_free(hp)
    e.close()
    return {"what": "n1: 2^25-frame 16-bit stereo WAV image (pinned host) -> HBM clip, zl_clip_create_from_wav incl. allocation, H2D and device decode; "
                    "n3: 2^24 stereo frames of host floats -> 16-bit WAV image, zl_wav_encode incl. H2D, device convert, D2H",
            "ingest_ms": round(ingest, 3), "ingest_file_GBps": nbytes / (ingest * 1e-3) / 1e9,
            "encode_ms": round(min(t_enc[1:]), 3), "encode_checksum": int(out[104:104 + 4096].astype(np.int64).sum())}


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus > 1 and world == 1 and args.impl == "ours":
        # convenience: relaunch under torchrun (the driver launches torchrun itself)
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29541", os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    workload = (f"cfg2x{args.sessions}: {args.sessions} independent 128-voice polyphonic sessions per GPU (BASELINE configs[1] "
                f"shape = 16 buses x 8 voices, batched as configs[4] offline render), {args.blocks} periods x {FRAMES} frames per step")
    config = {"workload": workload, "sessions_per_gpu": args.sessions, "buses": BUSES, "voices_per_bus": VOICES,
              "frames": FRAMES, "blocks_per_step": args.blocks, "mode": "Z (libzl process(), quirks faithful)",
              "outputs": "raw bus + master + peak meters" + (" + RMS sums" if args.rms else " (RMS, an extension over libzl's peak-only AudioLevels: --rms)"), "quirks": args.quirks,
              "kernels_per_period": "plan_kernel + walk_kernel + gather_kernel (+ one master_mix_kernel per step)", "parallelism": (f"sessions sharded over {args.gpus} GPU(s), no collective" if args.span_mix <= 1 else
                              f"sessions sharded over {args.gpus} GPU(s); groups of {args.span_mix} GPUs share one mix: zl_mix_reduce (ncclReduce of the master planes to the group's first rank) per step, the root alone reads the mix back"),
              "l2": "inputs (clip PCM 3 MB/session) larger than L2; no flush needed"}

    if args.impl == "reference":
        if rank != 0:
            return
        sessions = cpu_sample_size(args)
        vf, dt, th = run_cpu(args, sessions, args.steps, min(args.warmup, 1))
        v = vf / dt
        # this arm renders a bounded SAMPLE of the workload (same per-session shape, fewer sessions): the line says so, and
        # ms_per_step is the sample's own; the throughput in voice-frames/s is what compares (it does not depend on the count)
        config = dict(config, sessions_per_gpu=sessions,
                      workload=f"cfg2x{sessions} (CPU sample of cfg2x{args.sessions}: {sessions} of the {args.sessions} sessions per GPU, same 128-voice "
                               f"session shape, {args.blocks} periods x {FRAMES} frames per step)",
                      full_workload_sessions_per_gpu=args.sessions, parallelism=f"{th} host threads, sessions split across them",
                      kernels_per_period="n/a (CPU)")
        line = {"impl": "reference", "metric": "voice_frames_per_s", "value": v, "unit": "voice-frames/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
                "ms_per_step_is_for": f"the {sessions}-session sample; the full {args.sessions}-session step would take {1e3 * dt / args.steps * args.sessions / sessions:.0f} ms at this rate",
                "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32 (f64 positions)", "data": "synthetic", "config": config,
                "cpu_baseline": {"value": v, "unit": "voice-frames/s", "cores": th, "kind": "port",
                                 "sample": f"{sessions} sessions x {args.blocks} periods x {FRAMES} frames per step, {args.steps} steps "
                                           "(oracle port built -O3 -mfma -ffp-contract=fast; libzl itself needs Qt5/JACK and cannot be built here)"},
                "e2e": {"value": v, "unit": "voice-frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line), flush=True)
        return

    import torch
    r = run_cuda(args, rank, world, local_rank)
    elapsed = torch.tensor([r["elapsed_ms"], r["e2e_s"] * 1e3, r["e2e_pcm_s"] * 1e3], dtype=torch.float64, device=r["dev"])
    counts = torch.tensor([r["timed_vf"], r["e2e_vf"], r["timed_src"], r["e2e_pcm_vf"]], dtype=torch.float64, device=r["dev"])
    if r["dist"] is not None:
        r["dist"].all_reduce(elapsed, op=r["dist"].ReduceOp.MAX)
        r["dist"].all_reduce(counts, op=r["dist"].ReduceOp.SUM)
    el_ms, e2e_ms, e2e_pcm_ms = elapsed.tolist()
    vf, e2e_vf, src, e2e_pcm_vf = counts.tolist()
    if rank == 0:
        peak, peak_src = peaks()
        # roofline of the dominant kernel (gather_kernel, one launch = one 512-frame period of every session on this GPU):
        # algorithmic bytes per launch = unique source samples spanned by the voices (counted by the planner) + the raw
        # bus planes + the integer peaks, all per period; duration = mean of the CUDA-event pairs around each launch
        step_alg_bytes = (4.0 * r["timed_src"] / args.steps) + r["out_bytes"] + r["state_bytes"]
        gather_bytes = (4.0 * r["timed_src"] / args.steps + S_BUS_BYTES(args) + args.blocks * args.sessions * BUSES * 2 * 4) / args.blocks
        alg_bytes = gather_bytes
        mean_ms = r["kernel_ms"] or r["gather_alone_ms"]          # --no-kernel-events: fall back to the stand-alone launch
        ach = alg_bytes / (mean_ms * 1e-3) / 1e9
        line = {"metric": "voice_frames_per_s", "value": vf / (el_ms * 1e-3), "unit": "voice-frames/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": el_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32 (f64 positions)", "data": "synthetic", "config": config,
                "clocks": r["clocks"], "gpu_launches": r["launches"],
                # end to end = the offline bounce a host makes: commands in (pinned host, H2D), every session's mix out as the
                # 16-bit PCM of the reference recorder's WAV files (lib/AudioLevels.cpp:35-119), on the host when the clock stops.
                # The same loop with the masters read back as f32 planes (the parity path) is reported beside it as e2e_f32:
                # at 8 GPUs behind one host it is bound by the host's memory/PCIe, not by anything on the devices.
                "e2e": {"value": e2e_pcm_vf / (e2e_pcm_ms * 1e-3), "unit": "voice-frames/s", "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h_pcm"],
                        "what": "per step: zl_command_batch (H2D from pinned host memory) + zl_render + zl_read_master_all_pcm_async(bits=16) - every session's master "
                                "leaves the device as the 16-bit data chunk of the reference recorder's WAV file, encoded on the device, D2H into pinned host memory "
                                "overlapping the next step's render; zl_read_wait before the clock stops; bytes checked against zl_bounce_wav afterwards"},
                "e2e_f32": {"value": e2e_vf / (e2e_ms * 1e-3), "unit": "voice-frames/s", "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"],
                            "what": "the same loop with zl_read_master_all_async: the masters as f32 planes (twice the D2H bytes; the path the parity_at_size check reads)"},
                "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                             "traffic": (ncu_traffic() if (args.sessions, args.blocks, args.quirks) == (4096, 16, 7) else None),
                             "traffic_source": f"{TRAFFIC_FILE} (ncu --set full, builder-run capture of this command; NOT measured by this run)",
                             "peak_source": peak_src, "frac_of_nominal_8TBs": ach / 8000.0,
                             "algorithmic_bytes_per_launch": alg_bytes,
                             "bytes_per_voice_frame": step_alg_bytes / (r["timed_vf"] / args.steps),
                             "kernel": "zl::gather_kernel", "kernel_ms": mean_ms, "launches_per_step": args.blocks,
                             "gather_ms_per_step": r["gather_ms_per_step"],
                             "note": "kernel_ms is measured live in the timed region, where plan_kernel/walk_kernel of the next period run beside gather_kernel on a second stream and share its SMs; kernel_alone_ms is the same launch with the overlap switched off (one extra untimed step)",
                             "kernel_alone_ms": r["gather_alone_ms"], "frac_alone": alg_bytes / (r["gather_alone_ms"] * 1e-3) / 1e9 / peak,
                             "whole_step_frac": step_alg_bytes / (el_ms / args.steps * 1e-3) / 1e9 / peak,
                             "whole_step_note": "all algorithmic bytes of a step (source + voice state + raw planes + master + peaks) / ms_per_step: plan, walk, gather and master mix together"},
                "voice_frames_per_step": vf / args.steps / world, "host_submit_ms_per_step": r["host_submit_ms"],
                "host_submit_note": "host time inside zl_render per step (commands, launches), without the wait for a free command slot; "
                                    f"the submitting loop itself took {r['host_loop_ms']:.2f} ms per step because it runs ahead of the device and is held back",
                "parity_at_size": r["parity"]}
        if world == 1 and not args.no_latency:
            try:
                line["latency"] = latency_probe(local_rank)
                line["cpu_latency"] = cpu_latency_probe()
                line["scheduler"] = sched_probe()
            except Exception as ex:   # noqa: BLE001
                line["latency"] = {"error": str(ex)}
        if world == 1 and not args.no_thumb:
            try:
                del r["keep"]
                line["thumbnail"] = thumbnail_probe(local_rank, peak)
            except Exception as ex:   # noqa: BLE001
                line["thumbnail"] = {"error": str(ex)}
            try:
                line["wav_io"] = wav_io_probe(local_rank)
            except Exception as ex:   # noqa: BLE001
                line["wav_io"] = {"error": str(ex)}
        if world == 1 and not args.no_cpu_baseline:
            sessions = cpu_sample_size(args)
            cvf, cdt, th = run_cpu(args, sessions, 2, 1)
            line["cpu_baseline"] = {"value": cvf / cdt, "unit": "voice-frames/s", "cores": th, "kind": "port",
                                    "sample": f"{sessions} sessions x {args.blocks} periods x {FRAMES} frames, 2 steps (oracle port, -O3 -mfma contraction on)"}
        print(json.dumps(line), flush=True)
    r["e"].close()
    if r["dist"] is not None:
        r["dist"].barrier()
        r["dist"].destroy_process_group()


if __name__ == "__main__":
    main()
