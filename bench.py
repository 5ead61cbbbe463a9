#!/usr/bin/env python
"""bench.py — the scene-update frame (transform + bounds + 11-view cull + draw-list compaction)
at 16M entities per GPU, BASELINE.json's metric.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--entities E]

One "step" = one frame of the hot path over the whole scene: pmb_frame (update_scene loops A-C,
then filter + frustum cull + compaction for 11 views).  `value` = entities/s with the component
arrays resident in HBM; `e2e` = the same frame through pmb_frame_host with pinned HOST buffers
(H2D of every input array, kernels, D2H of counts + draw lists + renderable extents inside the
timed region).  N > 1: one process per GPU, every rank owns a root-aligned shard of 16M entities
(weak scaling, config 5 = 128M at N = 8) of ONE scene cut by synth.partition_by_root; the counts table and the
renderable extents go to every rank over NVLink by the library's own peer-store kernels and every GPU keeps the draw
lists of its shard with --exchange counts (each GPU draws what it culled); the default, `root`, gathers the lists on
rank 0 (BASELINE config 5) and times the counts-only mode beside it; `all` replicates them, `nccl` is the all-gather
fallback.  Outside the timed region the N ranks also run a 1 Mi-entity-per-rank scene through the same exchange and
rank 0 compares the gathered lists with the oracle's of the unsharded scene (config.exchange.oracle_parity).

--impl reference times the reference's CPU path (the oracle's restatement of it: the reference
itself does not compile here) on the host cores on a bounded sample of the workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "entities/sec per frame (transform+bounds+cull) at 16M"
UNIT = "entities/s"
BYTES_PER_ENTITY = 232  # 108 B read + 124 B written per entity for stages 1-2 (+ cull fused)
NUM_VIEWS = 11


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """SM clock + throttle reasons sampled DURING the timed region (B200_PROFILING.md): NVML
    (pynvml, what nvidia-smi reads) every 5 ms; one nvidia-smi query as the fallback."""

    def __init__(self, index):
        self.index = index
        self.sm, self.mx, self.power, self.when = [], [], [], []
        self.reasons = set()
        self._stop = threading.Event()
        self._t = None
        self._h = None
        self._nv = None
        try:
            import pynvml
            pynvml.nvmlInit()
            # NVML enumerates physical order; match CUDA's device by PCI bus id
            import torch
            bus = torch.cuda.get_device_properties(index).pci_bus_id
            dom = torch.cuda.get_device_properties(index).pci_domain_id
            dev = torch.cuda.get_device_properties(index).pci_device_id
            self._h = pynvml.nvmlDeviceGetHandleByPciBusId(f"{dom:08x}:{bus:02x}:{dev:02x}.0")
            self._nv = pynvml
        except Exception:
            self._h = None

    def _sample_nvml(self):
        nv, h = self._nv, self._h
        self.when.append(time.perf_counter())
        self.sm.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
        self.mx.append(float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)))
        try:
            self.power.append(nv.nvmlDeviceGetPowerUsage(h) / 1000.0)
        except Exception:
            pass
        r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
        for name, bit in (("hw_slowdown", nv.nvmlClocksEventReasonHwSlowdown),
                          ("hw_thermal_slowdown", nv.nvmlClocksEventReasonHwThermalSlowdown),
                          ("sw_thermal_slowdown", nv.nvmlClocksEventReasonSwThermalSlowdown),
                          ("sw_power_cap", nv.nvmlClocksEventReasonSwPowerCap)):
            if r & bit:
                self.reasons.add(name)

    def _sample_smi(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                             capture_output=True, text=True, timeout=5).stdout
        parts = [p.strip() for p in out.strip().split(",")]
        self.when.append(time.perf_counter())
        self.sm.append(float(parts[0]))
        self.mx.append(float(parts[1]))
        for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[3:7]):
            if val.lower().startswith("active"):
                self.reasons.add(name)

    def _run(self):
        while not self._stop.is_set():
            try:
                if self._h is not None:
                    self._sample_nvml()
                else:
                    self._sample_smi()
            except Exception:
                pass
            self._stop.wait(0.001)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self, t0=None, t1=None):
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        out = {"sm_mhz": float(np.median(self.sm)), "sm_max_mhz": float(max(self.mx)), "reasons": sorted(self.reasons),
               "samples": len(self.sm), "source": "nvml" if self._h is not None else "nvidia-smi"}
        if t0 is not None:
            out["samples_in_timed_region"] = int(sum(1 for w in self.when if t0 <= w <= t1))
            out["sampled"] = "from the warm-up to the end of a tail of further untimed steps of the same load (a query takes longer than a step)"
        if self.power:
            out["power_w_max"] = float(max(self.power))
        return out


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_frame_rate(n_sample, seed, omp, min_seconds, max_frames=20):
    """entities/s of the reference CPU path (oracle restatement) on an n_sample-entity scene of
    the bench distribution with the same 11 views."""
    from oracle import oracle as orc
    from pmtech_b200 import synth
    sc = synth.city_scene(n_sample, seed=seed)
    views = synth.city_views(orc.OracleCamera)
    o = orc.OracleScene(sc["parents"], sc["local_matrices"], sc["entities"], sc["state_flags"], sc["bounding_volumes"])
    if omp:
        orc.set_num_threads()
    o.frame(views, omp=omp)  # warm: page in, spin the thread pool up
    frames, threads, t0 = 0, 1, time.perf_counter()
    times = []
    while frames < max_frames and (frames < 2 or time.perf_counter() - t0 < min_seconds):
        a = time.perf_counter()
        _, threads = o.frame(views, omp=omp)
        times.append(time.perf_counter() - a)
        frames += 1
    return n_sample / float(np.median(times)), threads, frames, float(np.median(times))


def run_reference(args, rank):
    if rank != 0:
        return
    n_sample = args.ref_entities
    steps, warm = max(args.steps, 1), max(args.warmup, 0)
    from oracle import oracle as orc
    from pmtech_b200 import synth
    sc = synth.city_scene(n_sample, seed=3)
    views = synth.city_views(orc.OracleCamera)
    o = orc.OracleScene(sc["parents"], sc["local_matrices"], sc["entities"], sc["state_flags"], sc["bounding_volumes"])
    # the reference as it is: one thread (ecs_scene.cpp:1405 runs on the user thread) — two frames are enough to quote it
    o.frame(views, omp=False)
    t0 = time.perf_counter()
    single_frames = 1 if n_sample > (1 << 22) else 3
    for _ in range(single_frames):
        o.frame(views, omp=False)
    single = n_sample / ((time.perf_counter() - t0) / single_frames)
    orc.set_num_threads()  # all host cores, also under torchrun (which exports OMP_NUM_THREADS=1)
    threads = 1
    for _ in range(warm):
        _, threads = o.frame(views, omp=True)
    t0 = time.perf_counter()
    for _ in range(steps):
        _, threads = o.frame(views, omp=True)
    dt = (time.perf_counter() - t0) / steps
    value = n_sample / dt
    same = n_sample == args.entities
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_text(n_sample),
                   "entities_per_gpu": n_sample, "entities_per_step": n_sample, "views": NUM_VIEWS,
                   "same_config": bool(same)},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                         "single_thread": {"value": single, "cores": 1, "frames": single_frames,
                                           "what": "the reference as written: update_scene and the per-view culls on one thread"},
                         "sample": f"the bench scene itself ({n_sample} entities, seed 3, {NUM_VIEWS} views), every step one whole "
                                   f"frame: update_scene loops A-C + per-view filter+cull, OpenMP over all host threads ({threads}) "
                                   "— more than the single-threaded reference uses; the reference itself does not compile "
                                   "here (third_party/maths absent)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


_JSON_OUT = None


def emit(line):
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def run_cpu_configs():
    """cpu_baseline leg for the BASELINE configs other than the headline one:
    the oracle's restatement of the reference loops, single thread and OpenMP, one JSON line each."""
    from oracle import oracle as orc
    from pmtech_b200 import synth
    cases = [("config1: 64k torus, 1 camera", synth.torus_scene(seed=1), synth.torus_views(orc.OracleCamera), "aabb"),
             ("config2: 1M entities, depth 16, update only", synth.deep_scene(), None, "aabb"),
             ("config4: 10k rigs x 64 bones, 6 views", synth.rig_scene(10_000, seed=4), synth.mixed_views(orc.OracleCamera), "sphere")]
    for name, sc, views, mode in cases:
        n = sc["num_entities"]
        out = {"config": name, "entities": n}
        for omp in (False, True):
            o = orc.OracleScene(sc["parents"], sc["local_matrices"], sc["entities"], sc["state_flags"], sc["bounding_volumes"])
            times, threads = [], 1
            for _ in range(5):
                t0 = time.perf_counter()
                if views is None:
                    o.update_scene()
                else:
                    _, threads = o.frame(views, mode=mode, omp=omp)
                times.append(time.perf_counter() - t0)
            if views is None and omp:
                continue  # the update-only entry point has no OpenMP variant
            key = "all_cores" if omp else "single_thread"
            out[key] = {"entities_per_s": n / float(np.median(times[1:])), "ms": float(np.median(times[1:])) * 1e3, "threads": threads}
        emit(out)


def shim_e2e(capi, np, sc, views, steps, pinned=True):
    """The frame through the C++ binding a pmtech build would ship (include/put_ecs_b200.hpp over the mock of pmtech's types,
    tests/shim/shim_driver.cpp): update_scene in its resident form (every transform dirtied: 40 B cmp_transform records up,
    flags set on the device, nothing read back but the extents), all views in one batched pass, then one
    render_scene_view-style list fetch per view into host memory.  Host wall clock: every call synchronises."""
    import ctypes as C
    shim_dir = os.path.join(ROOT, "tests", "shim")
    so = os.path.join(shim_dir, "libshim_driver.so")
    src = os.path.join(shim_dir, "shim_driver.cpp")
    deps = [src, os.path.join(shim_dir, "mock_pmtech.h"), os.path.join(ROOT, "include", "put_ecs_b200.hpp"),
            os.path.join(ROOT, "include", "pmtech_b200.h")]
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(d) for d in deps):
        libdir = os.path.dirname(capi.SO_PATH)
        subprocess.check_call(["g++", "-O2", "-std=c++11", "-fPIC", "-shared", "-I" + shim_dir, "-o", so, src, "-L" + libdir,
                               "-lpmtech_b200", "-Wl,-rpath," + libdir])
    L = C.CDLL(so)
    L.shim_view_list.restype = C.c_int
    n, nv = sc["num_entities"], len(views)
    alloc = (lambda shape, dt: capi.pinned_empty(shape, dt)) if pinned else (lambda shape, dt: np.empty(shape, dt))
    arr = {"entities": alloc((n,), np.uint64), "state_flags": alloc((n,), np.uint64), "parents": alloc((n,), np.uint32),
           "local": alloc((n, 16), np.float32), "world": alloc((n, 16), np.float32), "bv": alloc((n, 13), np.float32),
           "pe": alloc((n, 8), np.float32)}
    has_tr = sc.get("transforms") is not None  # else the 64 B local matrices travel
    if has_tr:
        arr["tr"] = alloc((n, 10), np.float32)
    try:
        arr["entities"][...] = sc["entities"] | np.uint64(1 << 12)
        arr["state_flags"][...] = sc["state_flags"]
        arr["parents"][...] = sc["parents"]
        arr["local"][...] = sc["local_matrices"]
        arr["bv"][...] = sc["bounding_volumes"]
        if has_tr:
            arr["tr"][...] = sc["transforms"]
        arr["world"][...] = 0
        arr["pe"][...] = 0
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        L.shim_bind(C.c_uint32(n), C.c_uint32(n), p(arr["entities"]), p(arr["state_flags"]), p(arr["parents"]), p(arr["local"]),
                    p(arr["world"]), p(arr["bv"]), p(arr["pe"]))
        ext = np.zeros(6, np.float32)
        ptr_tr = p(arr["tr"]) if has_tr else None
        L.shim_update_opts(p(ext), None, 1, None, C.c_uint32(0), ptr_tr, 0, 0)  # first frame: everything goes up once
        fr = np.ascontiguousarray(views, np.float32)
        frs = [np.ascontiguousarray(v, np.float32) for v in views]
        out = capi.pinned_empty((n,), np.uint32) if pinned else np.empty(n, np.uint32)
        dirty = np.array([[0, n]], np.uint32)
        counts = [0] * nv

        def frame():
            L.shim_update_opts(p(ext), None, 0, p(dirty), C.c_uint32(1), ptr_tr, 1, 0)
            L.shim_prepare_views(p(fr), C.c_uint32(nv), 0)
            for v in range(nv):
                counts[v] = L.shim_view_list(p(frs[v]), 0, p(out), n)
        for _ in range(2):
            frame()
        t0 = time.perf_counter()
        for _ in range(steps):
            frame()
        ms = (time.perf_counter() - t0) * 1e3 / steps
        L.shim_release()
        if pinned:
            capi.pinned_free(out)
        return {"value": n / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "entities": n, "views": nv,
                "h2d_bytes_per_step": int(arr["tr"].nbytes if has_tr else arr["local"].nbytes), "d2h_bytes_per_step": int(sum(counts)) * 4 + 24 + nv * 4,
                "visible_per_view": [int(c) for c in counts],
                "what": "b200::update_scene(frame_options: dirty range [0, n), cmp_transform records, e_cmp::transform set on the "
                        "device, no read-back) + b200::prepare_views (all views, one device pass) + one b200::culled_entities per "
                        "view (stretchy-buffer copy of its list), host arrays " + ("page-locked" if pinned else "pageable")
                        + ("" if has_tr else "; this scene carries no cmp_transform records: the 64 B local matrices travel")}
    finally:
        if pinned:
            for a in arr.values():
                capi.pinned_free(a)


def config1_e2e(capi, np, synth, Camera, DeviceScene, shim_e2e):
    """BASELINE config 1 — the scene pmtech's own demo runs (65,544 entities, one camera; readme.md:56,
    examples/code/entities/entities.cpp:131-184) — end to end with host buffers: through pmb_frame_host and through the C++
    binding, next to the CPU restatement's time for the same frame (single thread, like the reference)."""
    from oracle import oracle as orc
    sc = synth.torus_scene(seed=1)
    n = sc["num_entities"]
    views = synth.torus_views(Camera)
    out = {"entities": n, "views": 1}
    with DeviceScene(n, 0) as dev:
        pin = {k: capi.pinned_empty(sc[k].shape, sc[k].dtype) for k in ("local_matrices", "entities")}
        pin["local_matrices"][...] = sc["local_matrices"]
        pin["entities"][...] = sc["entities"] | np.uint64(1 << 12)
        for a, k in ((capi.PARENTS, "parents"), (capi.STATE_FLAGS, "state_flags"), (capi.BOUNDING_VOLUMES, "bounding_volumes")):
            dev.upload(a, sc[k])
        dev.synchronize()
        lists = [capi.pinned_empty((n,), np.uint32)]
        cnt = capi.pinned_empty((1,), np.uint32)
        ext = capi.pinned_empty((6,), np.float32)

        def frame():
            dev.frame_host(n, views, local_matrices=pin["local_matrices"], entities=pin["entities"],
                           renderable_extents_out=ext, lists_out=lists, counts_out=cnt)
        for _ in range(5):
            frame()
        t0 = time.perf_counter()
        steps = 200
        for _ in range(steps):
            frame()
        ms = (time.perf_counter() - t0) * 1e3 / steps
        out["frame_host"] = {"value": n / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms,
                             "h2d_bytes_per_step": int(pin["local_matrices"].nbytes + pin["entities"].nbytes),
                             "d2h_bytes_per_step": int(cnt[0]) * 4 + 28, "visible": int(cnt[0]),
                             "what": "pmb_frame_host: every local matrix (64 B) and entity word (8 B) re-sent from page-locked memory, "
                                     "frame in one launch, list + extents read back"}
        for a in list(pin.values()) + lists + [cnt, ext]:
            capi.pinned_free(a)
    o = orc.OracleScene(sc["parents"], sc["local_matrices"], sc["entities"], sc["state_flags"], sc["bounding_volumes"])
    ov = synth.torus_views(orc.OracleCamera)
    o.frame(ov)
    times = []
    for _ in range(10):
        o.entities[:] = sc["entities"] | np.uint64(1 << 12)
        t0 = time.perf_counter()
        o.frame(ov)
        times.append(time.perf_counter() - t0)
    try:
        out["shim"] = shim_e2e(capi, np, sc, views, 100)
    except Exception as ex:  # noqa: BLE001
        out["shim"] = {"unavailable": f"{type(ex).__name__}: {ex}"}
    out["cpu_single_thread"] = {"value": n / float(np.median(times)), "unit": UNIT, "ms_per_step": float(np.median(times)) * 1e3,
                                "kind": "port", "cores": 1}
    return out


def check_exchange(torch, dist, dev, gather, kind, counts, base_index, nv, rank, world, device="cuda", local_list=None):
    """Once, outside the timed region: what the last exchange left is what the protocol promises.  Every rank: the counts
    table holds every rank's own list lengths, the view totals are their sums, the extents are the min / max over the
    shards.  A rank that holds the global lists: view v = the ranks' segments in rank order, each segment = that rank's
    local list + its base index — checked on the first and last entries of every segment and on the whole of this rank's
    own segments; ascending across segment borders."""
    if local_list is None:
        from pmtech_b200 import shard

        def local_list(v, c):
            return torch.as_tensor(shard._DeviceU32(dev.device_list_ptr(v), c), device=device)
    mine = torch.tensor([int(c) for c in counts], dtype=torch.int64, device=device)
    table = torch.empty(world * nv, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(table, mine)
    table = table.view(world, nv).cpu().numpy()
    ext = torch.tensor(dev.renderable_extents(), device=device)
    mn, mx = ext[:3].clone(), ext[3:].clone()
    dist.all_reduce(mn, op=dist.ReduceOp.MIN)
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    info = {"kind": kind, "global_visible_per_view": [int(x) for x in table.sum(axis=0)]}
    ok = True
    if kind.startswith("peer"):
        r = gather.result
        ok &= bool(np.array_equal(r["counts"], table)) and bool(np.array_equal(r["view_totals"], table.sum(axis=0)))
        ok &= bool(np.array_equal(r["extents"].view(np.uint32), torch.cat([mn, mx]).cpu().numpy().view(np.uint32)))
        info["exchange_device_ms"] = r["device_ms"]
        info["renderable_extents"] = [float(x) for x in r["extents"]]
    holds = kind == "nccl" or gather.receives
    if holds:
        offs = np.concatenate([np.zeros((1, nv), np.int64), np.cumsum(table, axis=0)])  # offs[r, v]: start of rank r's segment
        for v in range(nv):
            g = gather.global_list(v)
            ok &= int(g.shape[0]) == int(table[:, v].sum())
            ends = []
            for q in range(world):
                c = int(table[q, v])
                if c:
                    a = int(offs[q, v])
                    ends += [int(g[a].item()) & 0xFFFFFFFF, int(g[a + c - 1].item()) & 0xFFFFFFFF]
                    lo, hi = q * dev.capacity, (q + 1) * dev.capacity  # equal shards: rank q's entities are [q n, (q + 1) n)
                    ok &= lo <= ends[-2] <= ends[-1] < hi
            ok &= all(a <= b for a, b in zip(ends, ends[1:]))
            c = int(counts[v])
            if c:  # this rank's own segment, entry for entry
                a = int(offs[rank, v])
                ok &= bool(torch.equal(g[a:a + c], local_list(v, c) + base_index))
    t = torch.tensor([1 if ok else 0], device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if not int(t.item()):
        raise SystemExit("bench.py: the exchange left lists / counts / extents that differ from what the shards hold")
    info["checked"] = "counts table, view totals, extents on every rank; segment ends, order and this rank's own segments on every rank that holds the lists"
    return info


def time_exchange_mode(torch, dist, dev, shard, stream, n, views, nv, base_index, mode, steps, warmup, world):
    """The same step (pmb_frame + exchange launch) with another exchange mode, timed like the headline: device events on
    the launching stream behind the last exchange, barrier on both sides, max over ranks.  Returns ms per step."""
    g, kind = shard.make_gather(dev, nv, base_index, n, scene_stream=stream, prefer="peer", mode=mode)

    def settle():
        g.finish()
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
    for _ in range(warmup):
        dev.frame(n, views)
        g.launch()
    settle()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        dev.frame(n, views)
        g.launch()
    g.wait_on(stream)
    e1.record(stream)
    settle()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if hasattr(g, "close"):
        g.close()
    return float(t.item()) / steps, kind


def oracle_parity(torch, dist, capi, shard, synth, Camera, DeviceScene, stream, local_rank, rank, world, n_p, mode):
    """Outside the timed region, where the driver runs N ranks: ONE scene of world x n_p entities, one root-aligned shard
    per rank through the same pmb_frame + exchange as the bench step; the rank that holds the gathered lists compares them,
    the counts table and the renderable extents with the oracle's update_scene + culls of the UNSHARDED scene (the oracle
    here is the checker, never the thing measured).  Three exchanges: both buffer sets and a reuse."""
    views = synth.city_views(Camera)
    nv = len(views)
    sc = synth.city_scene(n_p, seed=103 + rank)
    ok, info = True, {"entities_per_rank": n_p, "entities": n_p * world, "mode": mode}
    with DeviceScene(n_p, local_rank) as d2:
        d2.set_stream(stream.cuda_stream)
        keep = [d2.upload(capi.PARENTS, sc["parents"]), d2.upload(capi.LOCAL_MATRICES, sc["local_matrices"]),
                d2.upload(capi.ENTITIES, sc["entities"]), d2.upload(capi.STATE_FLAGS, sc["state_flags"]),
                d2.upload(capi.BOUNDING_VOLUMES, sc["bounding_volumes"])]
        d2.synchronize()
        del keep
        d2.frame(n_p, views)
        g, kind = shard.make_gather(d2, nv, rank * n_p, n_p, scene_stream=stream, prefer="peer", mode=mode)
        for _ in range(3):
            d2.frame(n_p, views)
            g.launch()
        total = g.finish()
        torch.cuda.synchronize()
        holds = kind == "nccl" or g.receives
        if rank == 0:
            from oracle import oracle as orc
            blocks = [sc] + [synth.city_scene(n_p, seed=103 + k) for k in range(1, world)]
            cat = {k: np.concatenate([b[k] for b in blocks]) for k in ("local_matrices", "entities", "state_flags", "bounding_volumes")}
            parents = np.concatenate([b["parents"] + np.uint32(k * n_p) for k, b in enumerate(blocks)])
            del blocks
            o = orc.OracleScene(parents, cat["local_matrices"], cat["entities"], cat["state_flags"], cat["bounding_volumes"])
            orc.set_num_threads()
            want, threads = o.frame(synth.city_views(orc.OracleCamera), omp=True)
            ok &= total == sum(len(l) for l in want)
            if kind.startswith("peer"):
                r = g.result
                ok &= bool(np.array_equal(r["counts"].sum(axis=0), [len(l) for l in want]))
                ok &= bool(np.array_equal(r["extents"].view(np.uint32), np.asarray(o.renderable_extents, np.float32).view(np.uint32)))
            if holds:
                for v in range(nv):
                    got = g.global_list(v).cpu().numpy().view(np.uint32)
                    ok &= bool(np.array_equal(got, want[v]))
            info.update({"oracle_threads": int(threads), "visible_per_view": [len(l) for l in want], "lists_compared": bool(holds),
                         "kind": kind})
        t = torch.tensor([1 if ok else 0], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        torch.cuda.synchronize()
        dist.barrier()
        if hasattr(g, "close"):
            g.close()
    info["oracle_parity"] = bool(int(t.item()))
    return info


def workload_text(n):
    """the `config.workload` of BOTH arms (the driver compares the two strings)"""
    return (f"{n}-entity scene per GPU, 90% roots / 10% depth 1-3 children, 1 perspective camera + 4 ortho cascades + "
            "6 point-light cube faces (11 views), transform + bounds + parent expansion + filter/cull + per-view draw lists")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--entities", type=int, default=1 << 24, help="entities per GPU")
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--cpu-entities", type=int, default=1 << 20, help="entities of the CPU-baseline sample")
    ap.add_argument("--ref-entities", type=int, default=1 << 24, help="entities per step of --impl reference (default: the bench scene)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--watchdog", type=float, default=480.0, help="seconds after which a stuck run gives up")
    ap.add_argument("--cpu-configs", action="store_true", help="time the CPU restatement on BASELINE configs 1, 2 and 4 and exit")
    ap.add_argument("--exchange", default="root", choices=["counts", "root", "all", "nccl"],
                    help="multi-GPU exchange, by peer stores: counts table + renderable extents to every rank and the draw lists "
                         "gathered on rank 0 (default: BASELINE config 5, 'NCCL-gathered visible lists') / left on the GPU that "
                         "draws them / replicated; nccl = all-gather of rows (replicates).  With the default the counts-only "
                         "mode is timed beside it (config.exchange_modes)")
    ap.add_argument("--parity-entities", type=int, default=1 << 20,
                    help="entities per rank of the multi-GPU oracle parity check that runs outside the timed region (0 = skip)")
    ap.add_argument("--shim-e2e", default=None, choices=["3", "1"], help=argparse.SUPPRESS)  # child mode of the e2e leg
    args = ap.parse_args()
    if args.shim_e2e:
        # runs in a process of its own: the binding asserts like the reference (PEN_ASSERT), and an assert must not take the bench line with it
        from pmtech_b200 import capi, synth
        from pmtech_b200.ecs import Camera
        from pmtech_b200.scene import DeviceScene
        if args.shim_e2e == "3":
            sc = synth.city_scene(args.entities, seed=3, with_transforms=True)
            res = shim_e2e(capi, np, sc, synth.city_views(Camera), max(args.e2e_steps, 3))
        else:
            res = config1_e2e(capi, np, synth, Camera, DeviceScene, shim_e2e)
        print(json.dumps(res), flush=True)
        return 0

    # NCCL and friends may write to fd 1 ("NCCL version ..."): keep the real stdout for the one JSON line
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    global _JSON_OUT
    _JSON_OUT = os.fdopen(json_fd, "w")
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.cpu_configs:
        if rank == 0:
            run_cpu_configs()
        return 0
    if args.impl == "reference":
        run_reference(args, rank)
        return 0
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if args.gpus > 1 and world == 1:
        raise SystemExit("for --gpus N > 1 launch with torch.distributed.run (one process per GPU)")
    args.warmup = max(args.warmup, 3)
    # a rank that never arrives must not hold the box until the caller's limit: leave loudly instead (torchrun then ends the others)
    watchdog = threading.Timer(args.watchdog, lambda: (sys.stderr.write(f"bench.py: rank {rank} made no progress for {args.watchdog} s, giving up\n"),
                                                       sys.stderr.flush(), os._exit(3)))
    watchdog.daemon = True
    watchdog.start()

    import torch
    import torch.distributed as dist
    from pmtech_b200 import capi, shard, synth
    from pmtech_b200.ecs import Camera
    from pmtech_b200.scene import DeviceScene

    if capi.device_count() < 1:
        raise SystemExit("bench.py needs a CUDA device: the scene-update path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    # ---- the scene: ONE scene of world x `entities`, cut at root boundaries into one shard per rank (weak scaling) ----
    # The global scene is `world` blocks of `entities`, block k generated from seed 3 + k; a block starts at a root and no
    # parent link leaves it.  Rank 0 runs the root-subtree partitioner over the global parents (config 5: 128 Mi entries)
    # and every rank takes its range from the cuts it returns.
    n = args.entities
    views = synth.city_views(Camera)
    nv = len(views)
    if world > 1:
        cuts = [None]
        if rank == 0:
            gp = np.concatenate([synth.city_parents(n, seed=3 + k) + np.uint32(k * n) for k in range(world)])
            cuts = [synth.partition_by_root(gp, world)]
            del gp
        dist.broadcast_object_list(cuts, src=0)
        first, last = cuts[0][rank]
        assert (first, last) == (rank * n, (rank + 1) * n), "the partitioner cuts the blocks where they were put together"
    else:
        first, last = 0, n
    sc = synth.city_scene(n, seed=3 + rank, with_transforms=not args.no_e2e)
    base_index = first  # global numbering of this shard's entities

    dev = DeviceScene(n, local_rank)
    stream = torch.cuda.Stream()  # a real stream: the handle, the CUDA events and the packs all work on this one
    torch.cuda.set_stream(stream)
    dev.set_stream(stream.cuda_stream)
    keep = [dev.upload(capi.PARENTS, sc["parents"]), dev.upload(capi.LOCAL_MATRICES, sc["local_matrices"]),
            dev.upload(capi.ENTITIES, sc["entities"]), dev.upload(capi.STATE_FLAGS, sc["state_flags"]),
            dev.upload(capi.BOUNDING_VOLUMES, sc["bounding_volumes"])]
    dev.synchronize()
    del keep

    gather, gather_kind = None, None
    if world > 1:
        dev.frame(n, views)  # the exchange sizes its buffers from a first frame's counts
        gather, gather_kind = shard.make_gather(dev, nv, base_index, n, scene_stream=stream,
                                                prefer="nccl" if args.exchange == "nccl" else "peer",
                                                mode=args.exchange if args.exchange != "nccl" else "root")

    def step():
        dev.frame(n, views)
        if gather is not None:
            gather.launch()  # pack on this stream, all-gather + unpack on the side stream: overlaps the next frame

    def barrier():
        if gather is not None:
            gather.finish()  # the last exchange has landed (every step's exchange is inside the timed region)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    clocks = ClockSampler(local_rank)  # from the warm-up on, every millisecond: the timed region alone is a few samples long
    clocks.__enter__()
    for _ in range(args.warmup):
        step()
    barrier()
    dev.reset_kernel_stats()
    dev.set_profiling(2)  # ONE event pair per step, around the dominant kernel: roofline.achieved is measured in the timed region
    launches0 = dev.total_launches()
    gather_launches0 = gather.launches if gather is not None else 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    wall0 = time.perf_counter()
    e0.record(stream)
    for _ in range(args.steps):
        step()
    if gather is not None:
        gather.wait_on(stream)  # the closing event is recorded behind the last exchange
    e1.record(stream)
    barrier()
    wall1 = time.perf_counter()
    ms_total = e0.elapsed_time(e1)
    launches = dev.total_launches() - launches0
    if gather is not None:
        launches += gather.launches - gather_launches0  # pack / publish / push (or pack / unpack beside NCCL's own)
    frame_timed = dev.kernel_stats()["frame"]  # the dominant kernel's launches of the timed region
    # the per-group breakdown (kernel_ms_by_group) comes from a SECOND pass of the same K steps, untimed: event pairs around
    # every launch group stay out of the headline's timed region
    dev.reset_kernel_stats()
    dev.set_profiling(1)
    for _ in range(args.steps):
        step()
    barrier()
    stats = dev.kernel_stats()
    dev.set_profiling(0)
    for _ in range(max(0, 400 - 2 * args.steps)):  # the same load a little longer (the same count on every rank): the clock
        step()                                     # sampler needs ~10 ms per query and the timed region may be shorter than that
    barrier()
    clocks.__exit__()
    counts = dev.counts(nv)
    exchange_info = None
    if gather is not None:
        exchange_info = check_exchange(torch, dist, dev, gather, gather_kind, counts, base_index, nv, rank, world)
        if args.parity_entities:
            par = oracle_parity(torch, dist, capi, shard, synth, Camera, DeviceScene, stream, local_rank, rank, world,
                                args.parity_entities, args.exchange if args.exchange != "nccl" else "root")
            exchange_info["oracle_parity"] = par["oracle_parity"]
            exchange_info["oracle_parity_check"] = par
            if not par["oracle_parity"]:
                raise SystemExit(f"bench.py: the {world}-rank exchange does not reproduce the oracle's lists of the unsharded scene: {par}")
    t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    total_entities = n * world
    value = total_entities / (ms_step * 1e-3)
    exchange_modes = None
    if gather is not None:
        exchange_modes = {args.exchange: {"ms_per_step": ms_step, "value": value, "kind": gather_kind}}
        if args.exchange == "root" and gather_kind == "peer-root":
            ms_c, kind_c = time_exchange_mode(torch, dist, dev, shard, stream, n, views, nv, base_index, "counts", args.steps,
                                              args.warmup, world)
            exchange_modes["counts"] = {"ms_per_step": ms_c, "value": total_entities / (ms_c * 1e-3), "kind": kind_c,
                                        "note": "counts table + extents only, the lists stay on the GPU that culled them"}

    # ---- roofline of the dominant kernel (frame_kernel: transform + bounds + plane tests) -------
    frame_ms, frame_launches = frame_timed
    peak, peak_src = peaks()
    k_ms = frame_ms / max(frame_launches, 1)
    alg_bytes = BYTES_PER_ENTITY * n
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9 if k_ms > 0 else 0.0
    traffic, traffic_commit = None, None
    tp = os.path.join(ROOT, "profiles", "frame_kernel_traffic.json")
    if os.path.exists(tp) and n == 1 << 24 and world == 1:  # an ncu figure of the 1-GPU run: not claimed for N > 1
        try:
            import hashlib
            rec = json.load(open(tp))
            here = hashlib.sha256(open(os.path.join(ROOT, "pmtech_b200", "csrc", "pm_kernels.cu"), "rb").read()).hexdigest()[:16]
            if rec.get("pm_kernels_sha16") == here:  # a capture of another kernel source is not this run's traffic
                traffic, traffic_commit = rec["dram_bytes_per_launch"], rec.get("commit")
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "kernel": "frame_kernel<MODE_AABB>", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic,
                "traffic_source": f"ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum of one launch at N = 1, captured "
                                  f"at commit {traffic_commit} on this pm_kernels.cu (profiles/frame_kernel_traffic.json)" if traffic else None,
                "peak_source": f"{peak_src} (MEASURED_PEAKS.json hbm_gbs)",
                "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": k_ms,
                "frac_of_nominal_8000": achieved / 8000.0,
                "step_gbs": (alg_bytes + 4.0 * float(counts.sum())) / (ms_step * 1e-3) / 1e9,
                "kernel_ms_source": "CUDA events around frame_kernel inside the timed region (pmb_set_profiling level 2: that kernel only)",
                "kernel_ms_by_group": {k: v[0] / args.steps for k, v in stats.items()},
                "kernel_ms_by_group_source": "a second, untimed pass of the same steps with events around every launch group"}

    # ---- e2e: host buffers in, draw lists out, through pmb_frame_host -----------------------------
    # The frame a pmtech demo produces (examples/code/entities/entities.cpp:211-216): the host dirtied
    # every transform, so this step's inputs are the 40 B cmp_transform records and the entity flags
    # (e_cmp::transform set); parents, bounds and state flags did not change and stay mirrored.  The
    # library bakes local = T*R*S on the device (stage 0), runs the frame, and the per-view counts,
    # draw lists and the renderable extents come back to host memory.  `full_upload` is the same call
    # with every input array re-sent (what a shim without dirty tracking would pay).
    e2e = None
    if not args.no_e2e:
        CMP_TRANSFORM = np.uint64(1 << 12)
        pin = {k: capi.pinned_empty(sc[k].shape, sc[k].dtype) for k in
               ("parents", "local_matrices", "entities", "state_flags", "bounding_volumes", "transforms")}
        for k in pin:
            pin[k][...] = sc[k]
        pin["entities"][...] = sc["entities"] | CMP_TRANSFORM
        lists_host = [capi.pinned_empty((int(c) + 4096,), np.uint32) for c in counts]
        counts_host = capi.pinned_empty((nv,), np.uint32)
        ext_host = capi.pinned_empty((6,), np.float32)

        def e2e_step(full=False, flags="device"):
            extra = dict(parents=pin["parents"], local_matrices=pin["local_matrices"], state_flags=pin["state_flags"],
                         bounding_volumes=pin["bounding_volumes"]) if full else {}
            if full or flags == "host":
                extra["entities"] = pin["entities"]          # e_cmp::transform travels inside the 8 B entity words
            else:
                dev.mark_transforms_dirty(0, n)              # ... or is set for the dirtied range on the device
            dev.frame_host(n, views, transforms=pin["transforms"], renderable_extents_out=ext_host,
                           lists_out=lists_host, counts_out=counts_host, **extra)
            if gather is not None:
                gather.run()  # host buffers in, global lists out: nothing to overlap with

        def timed(steps, full, flags="device"):
            for _ in range(2):
                e2e_step(full, flags)
            barrier()
            t0 = time.perf_counter()
            e0.record(stream)
            for _ in range(steps):
                e2e_step(full, flags)
            e1.record(stream)
            barrier()
            wall = (time.perf_counter() - t0) * 1e3
            ms = max(e0.elapsed_time(e1), wall)  # frame_host synchronises: host wall time is the honest clock
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item()) / steps

        # the ceiling of this step on this box: the bare pinned host->device copy of the same 40 B records, all ranks at once
        def h2d_ceiling():
            best = 0.0
            for _ in range(3):
                barrier()
                e0.record(stream)
                dev.upload(capi.TRANSFORMS, pin["transforms"])
                e1.record(stream)
                torch.cuda.synchronize()
                best = max(best, pin["transforms"].nbytes / (e0.elapsed_time(e1) * 1e-3) / 1e9)
            t = torch.tensor([best], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.SUM)
            return float(t.item())

        ceiling_gbs = h2d_ceiling()
        ms_e2e_step = timed(args.e2e_steps, False)
        ms_flags = timed(max(args.e2e_steps // 2, 2), False, "host")
        ms_full = timed(max(args.e2e_steps // 2, 2), True)
        h2d = int(pin["transforms"].nbytes)
        h2d_full = sum(int(pin[k].nbytes) for k in pin)
        d2h = int(counts_host[:nv].sum()) * 4 + nv * 4 + 24
        e2e = {"value": total_entities / (ms_e2e_step * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e_step,
               "pcie_ceiling_gbs": ceiling_gbs, "h2d_gbs": world * h2d / (ms_e2e_step * 1e-3) / 1e9,
               "frac_of_pcie_ceiling": (world * h2d / (ms_e2e_step * 1e-3) / 1e9) / ceiling_gbs if ceiling_gbs else None,
               "pcie_ceiling_what": "bare cudaMemcpyAsync of the same pinned 40 B transform records (671 MB per rank), every rank at "
                                    "once, best of 3, summed over the ranks; h2d_gbs = the same bytes over the e2e step",
               "what": "pmb_mark_transforms_dirty(0, n) + pmb_frame_host, every transform dirtied by the host: pinned "
                       "cmp_transform records (40 B) uploaded every step, e_cmp::transform set for the range on the device, "
                       "stage-0 bake + frame kernels, counts + 11 draw lists + renderable extents read back; parents / bounds "
                       "/ state flags stay mirrored",
               "flags_uploaded": {"value": total_entities / (ms_flags * 1e-3), "ms_per_step": ms_flags,
                                  "h2d_bytes_per_step": h2d + int(pin["entities"].nbytes),
                                  "what": "the same frame with the 8 B entity words (flag set by the host) uploaded too"},
               "full_upload": {"value": total_entities / (ms_full * 1e-3), "ms_per_step": ms_full, "h2d_bytes_per_step": h2d_full,
                               "what": "the same call with parents, local matrices, state flags and 52 B bound records re-sent too"}}
        for a in list(pin.values()) + lists_host + [counts_host, ext_host]:
            capi.pinned_free(a)
        # the same frame through the reference-side C++ binding, and pmtech's own operating point (config 1, the
        # 64 Ki-entity demo) both ways; N = 1 only.  Never fatal: the headline line does not depend on it.
        if world == 1:
            def child(which):
                try:
                    r = subprocess.run([sys.executable, os.path.abspath(__file__), "--shim-e2e", which, "--entities", str(n),
                                        "--e2e-steps", str(args.e2e_steps)], capture_output=True, text=True, timeout=300)
                    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
                    if r.returncode != 0 or not lines:
                        return {"unavailable": f"child exited {r.returncode}: {r.stderr.strip()[-300:]}"}
                    return json.loads(lines[-1])
                except Exception as ex:  # noqa: BLE001
                    return {"unavailable": f"{type(ex).__name__}: {ex}"}
            e2e["shim"] = child("3")
            if "value" in e2e["shim"]:
                e2e["shim"]["vs_frame_host"] = e2e["shim"]["value"] / e2e["value"]
            e2e["config1"] = child("1")

    # ---- CPU baseline on rank 0 (N = 1 only) -------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        v1, _, f1, s1 = cpu_frame_rate(args.cpu_entities, 3, omp=False, min_seconds=6.0)
        vN, thr, fN, sN = cpu_frame_rate(args.cpu_entities * 4, 3, omp=True, min_seconds=6.0)
        cpu = {"value": v1, "unit": UNIT, "cores": 1, "kind": "port",
               "sample": f"{args.cpu_entities}-entity scene of the bench distribution, {nv} views, {f1} frames "
                         f"(median {s1 * 1e3:.1f} ms): update_scene loops A-C + per-view filter + cull with sb_push lists, "
                         "single thread like the reference",
               "all_cores": {"value": vN, "cores": thr, "host_cores": host_cores(),
                             "sample": f"{args.cpu_entities * 4} entities, {fN} frames (median {sN * 1e3:.1f} ms), OpenMP "
                                       "variant of the same loops"}}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": workload_text(n), "total_entities": total_entities,
                       "entities_per_gpu": n, "views": nv, "visible_per_view": [int(c) for c in counts],
                       "partition": "one scene of all ranks' entities cut by partition_by_root (root subtree ranges), one shard per GPU" if world > 1 else "single GPU",
                       "global_visible_per_view": exchange_info["global_visible_per_view"] if exchange_info else None,
                       "gather": (({"peer-root": "counts table + renderable extents to every rank, draw lists gathered on rank 0: "
                                                 "NVLink peer stores from the library's kernels (16-byte stores into the final "
                                                 "per-view global lists), device-flag barrier, on a side stream",
                                    "peer-all": "the same with the lists replicated on every rank",
                                    "peer-counts": "counts table + renderable extents only; the lists stay sharded"}.get(
                                       gather_kind, "NCCL all_gather of fixed-size rows {counts, packed global-index lists} on a "
                                                    "side stream, device-side unpack (lists replicated)"))
                                  + f"; row_stride {gather.row_stride} u32, overflows {gather.overflows}") if world > 1 else None,
                       "exchange": exchange_info, "exchange_modes": exchange_modes,
                       "l2": "inputs larger than L2 (3.9 GB of component arrays per frame vs 126 MB L2)",
                       "fma": "no contraction (explicit __fmul_rn/__fadd_rn; oracle built -ffp-contract=off)"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
            "clocks": clocks.summary(wall0, wall1),
        }
        emit(line)
    watchdog.cancel()
    dev.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
