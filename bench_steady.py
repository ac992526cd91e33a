"""Steady-state reconcile sweeps for bench.py: the drift check driven the way the controller drives it.

The reference's caches are resident (informers keep them current, controller.go:286-359) and every 30 s
resync sweeps everything (main.go:70-71).  One *step* here is one such sweep after an interval's worth of
informer events:

    controller cluster   a fraction `churn` of its objects got a new value (one byte changed, same length)
    every shard          received the controller's values of the PREVIOUS interval (what UpdateSecret /
                         UpdateConfigMap / UpdateTemplate caused), so drift does not pile up: each sweep finds
                         about churn x objects x shards decisions != SKIP

GPU arm  (SteadyGpu):  ncdh_apply_events (canonicalise + cache update) -> ncdh_scan (pack changed records into
         pinned arenas, pull them over PCIe, scan, diff list back) -> ncdh_sweep_queue + a handful of
         ncdh_lookup — the DriftChecker API of INTEGRATION.md, libncdhost.so over libncdrift.so.
CPU arm  (SteadyCpu):  the same event records -> oracle fleet: parse the values, store them, sweep every template
         and workgroup over every shard with map lookups and reflect.DeepEqual restated on the parsed values.

Events are prebuilt (ncdh_batch_from_arrays) before a step is timed: they are the step's INPUT, host buffers
holding Go values in transport form, as an informer would hand over decoded objects.
"""
from __future__ import annotations

import ctypes as C
import os
import time

import numpy as np


def mem_available_gb() -> float:
    try:
        with open("/proc/meminfo") as f:
            for line in f:
                if line.startswith("MemAvailable:"):
                    return int(line.split()[1]) / 1e6
    except OSError:
        pass
    return 64.0


def converged_cfg(make_cfg, workload, groups, part_mod=0, part_rem=0):
    """The workload's shapes with no initial drift and the same object order in every cluster: the drift of
    the steady-state legs comes from the events alone, and object i is the same object everywhere."""
    cfg, S = make_cfg(workload, groups=groups, part_mod=part_mod, part_rem=part_rem)
    cfg.p_flip_ppm = cfg.p_missing_ppm = cfg.p_len_ppm = cfg.p_owner_ppm = cfg.p_rogue_ppm = 0
    cfg.shuffle = 0
    return cfg, S


def download_clusters(device, cfg, S):
    """Host images (HostArena) of the controller and the S shards from the on-device generator.  Shard images
    drop their payload: shard events take their VALUES from the controller (the shard received them)."""
    from nexus_configuration_controller_b200.drift import Context

    hosts = []
    with Context(device) as ctx:
        for role in range(-1, S):
            a = ctx.synth_arena(cfg, role)
            hosts.append(a.download())
            a.close()
    ctrl = hosts[0]
    for h in hosts[1:]:
        assert h.nobj == ctrl.nobj and np.array_equal(h.kind[: h.nobj], ctrl.kind[: ctrl.nobj]) and \
            np.array_equal(h.key_len[: h.nobj], ctrl.key_len[: ctrl.nobj]), "clusters are not laid out alike"
        for i in (0, h.nobj // 2, h.nobj - 1):
            assert h.key(i) == ctrl.key(i)
    return hosts


class Steps:
    """The event batches of the steps of one churn level (prebuilt: they are the steps' input)."""

    def __init__(self, hosts, churn, nsteps, seed, first_salt, prev_idx=None, prev_salt=0, threads=0, lib=None):
        from nexus_configuration_controller_b200.host import Event, EventBatch

        self._owners = []
        rng = np.random.default_rng(seed)
        ctrl, n = hosts[0], hosts[0].nobj
        k = max(1, int(round(n * churn)))
        self.steps, self.value_bytes, self.events = [], 0, 0
        idx_prev, salt_prev = prev_idx, prev_salt
        for s in range(nsteps):
            idx = np.sort(rng.choice(n, size=k, replace=False)).astype(np.uint32) if k < n else np.arange(n, dtype=np.uint32)
            salt = first_salt + s
            batches = [EventBatch(ctrl, cluster=-1, obj_idx=idx, salt=salt, threads=threads, lib=lib)]
            if idx_prev is not None:  # the shards catch up with the previous interval's controller values
                batches += [EventBatch(h, values=ctrl, cluster=c, obj_idx=idx_prev, salt=salt_prev, threads=threads, lib=lib)
                            for c, h in enumerate(hosts[1:])]
            # one array per step: an interval's events of every cluster go through ONE ncdh_apply_events call
            total = sum(b.n for b in batches)
            merged = (Event * max(total, 1))()
            at = 0
            for b in batches:
                C.memmove(C.byref(merged, at * C.sizeof(Event)), b.events, b.n * C.sizeof(Event))
                at += b.n
            self.steps.append((merged, total))
            self._owners.append(batches)
            idx_prev, salt_prev = idx, salt
        self.last_idx, self.last_salt = idx_prev, salt_prev
        self.value_bytes = sum(b.value_bytes for b in self._owners[-1])
        self.events = self.steps[-1][1]

    def close(self):
        for batches in self._owners:
            for b in batches:
                b.close()
        self.steps, self._owners = [], []


class SteadyGpu:
    """The fleet resident behind a DriftChecker (libncdhost.so)."""

    def __init__(self, hosts, device, threads=0, lib=None):
        from nexus_configuration_controller_b200.host import DriftChecker, EventBatch

        self.S = len(hosts) - 1
        self.f = DriftChecker(self.S, device=device, lib=lib)
        self.threads = threads
        t0 = time.time()
        for c, h in enumerate(hosts):  # initial list: every object of every cluster, controller values everywhere
            b = EventBatch(h, values=hosts[0], cluster=c - 1, threads=threads, lib=lib)
            self.f.apply_events(b, threads)
            b.close()
        self.f.scan()  # the first scan uploads everything
        self.load_s = time.time() - t0
        self.first = self.f.timing()
        self.ndiff0, self.decisions = self.f.counts()
        ctrl = hosts[0]
        heads = [i for i in range(min(ctrl.nobj, 4096)) if int(ctrl.kind[i]) in (0, 1)][:64]
        self.lookup_keys = [(int(ctrl.kind[i]),) + tuple(x.encode() for x in ctrl.key(i).split("/", 1)) for i in heads]
        self.qbuf = C.create_string_buffer(1 << 16)

    def step(self, batches):
        """One timed step.  Returns (seconds, timing dict of the scan)."""
        f, lib = self.f, self.f.lib
        events, n = batches
        t0 = time.perf_counter()
        if lib.ncdh_apply_events(f.h, events, n, self.threads) != 0:
            raise RuntimeError(lib.ncdh_last_error().decode())
        f.scan()
        n = lib.ncdh_sweep_queue(f.h, self.qbuf, len(self.qbuf))  # the work-queue feed
        if n > len(self.qbuf):
            self.qbuf = C.create_string_buffer(int(n * 2))
            lib.ncdh_sweep_queue(f.h, self.qbuf, len(self.qbuf))
        code = C.c_uint32()
        for kind, ns, name in self.lookup_keys:  # what handlers ask for: (template, dependent, shard)
            lib.ncdh_lookup(f.h, kind, ns, name, 0, 0, C.byref(code))
        dt = time.perf_counter() - t0
        return dt, f.timing(), n

    def close(self):
        self.f.close()


class SteadyCpu:
    """The same fleet as parsed objects behind the oracle's handlers (the CPU arm)."""

    def __init__(self, hosts, threads, lib=None):
        import oracle
        from nexus_configuration_controller_b200.host import EventBatch

        self.S, self.threads = len(hosts) - 1, threads
        self.fleet = oracle.Fleet(self.S)
        for c, h in enumerate(hosts):
            b = EventBatch(h, values=hosts[0], cluster=c - 1, threads=threads, lib=lib)
            self.fleet.apply(b.events, b.n, threads)
            b.close()
        self.last = self.fleet.sweep(threads)  # warm-up sweep

    def step(self, batches):
        events, n = batches
        t0 = time.perf_counter()
        self.fleet.apply(events, n, self.threads)
        self.last = self.fleet.sweep(self.threads)
        return time.perf_counter() - t0

    def close(self):
        self.fleet.close()


def median(xs):
    xs = sorted(xs)
    return xs[len(xs) // 2]


def run_levels(arm_gpu, arm_cpu, hosts, levels, steps_for, threads, lib=None):
    """Runs every churn level on the arms given (either may be None).  Returns {churn: row}."""
    out = {}
    prev_idx, prev_salt, salt = None, 0, 1
    for churn in levels:
        nsteps, warm = steps_for(churn)
        st = Steps(hosts, churn, nsteps + warm, seed=int(churn * 1e6) + 17, first_salt=salt, prev_idx=prev_idx, prev_salt=prev_salt,
                   threads=threads, lib=lib)
        salt += nsteps + warm
        gpu_t, cpu_t, tims, same, feed = [], [], [], True, 0
        for k, batches in enumerate(st.steps):
            if arm_gpu is not None:
                dt, tm, feed = arm_gpu.step(batches)
                if k >= warm:
                    gpu_t.append(dt)
                    tims.append(tm)
            if arm_cpu is not None:
                dt = arm_cpu.step(batches)
                if k >= warm:
                    cpu_t.append(dt)
            if arm_gpu is not None and arm_cpu is not None:
                _, dec, nd, dirty, dig = arm_cpu.last
                same = same and arm_gpu.f.counts() == (nd, dec) and arm_gpu.f.diff_digest() == (dig, dirty)
        prev_idx, prev_salt = st.last_idx, st.last_salt
        row = {"churn": churn, "events_per_step": st.events, "value_bytes_per_step": st.value_bytes, "steps": nsteps}
        if arm_gpu is not None:
            ms = median(gpu_t) * 1e3
            mid = tims[len(tims) // 2]
            row.update({
                "ms_per_step": ms, "decisions_per_s": arm_gpu.decisions / (ms / 1e3),
                "split_ms": {k: round(mid[k], 3) for k in ("canon_ms", "pack_ms", "upload_ms", "scan_ms", "gpu_scan_ms", "publish_ms")},
                "h2d_bytes_per_step": int(mid["h2d_bytes"]), "d2h_bytes_per_step": int(mid["d2h_bytes"]) + int(feed),
                "records_synced": int(mid["records_synced"]), "diff_entries": arm_gpu.f.counts()[0],
            })
            row["split_ms"]["other_ms"] = round(ms - mid["canon_ms"] - mid["total_ms"], 3)  # sweep queue, lookups, call overhead
        if arm_cpu is not None:
            ms = median(cpu_t) * 1e3
            row.update({"cpu_ms_per_step": ms, "cpu_decisions_per_s": arm_cpu.last[1] / (ms / 1e3), "cpu_diff_entries": arm_cpu.last[2]})
        if arm_gpu is not None and arm_cpu is not None:
            row["identical_to_cpu"] = bool(same)
        out[churn] = row
        st.close()
    return out


def default_steps(churn):
    """(timed steps, warm-up steps) per churn level: heavy levels get fewer."""
    if churn >= 0.5:
        return 1, 1
    if churn >= 0.05:
        return 3, 1
    return 7, 2


def break_even(ratios):
    """The churn at which the GPU arm's decisions/s falls to the CPU arm's AT THE SAME CHURN (both arms slow down
    as more events arrive).  ratios: {churn: gpu / cpu}.  Log-linear interpolation between measured levels, or a
    statement that the arms do not cross within the measured range."""
    import math

    pts = sorted((float(c), float(v)) for c, v in ratios.items())
    if all(v > 1 for _, v in pts):
        return {"churn": None, "note": f"GPU arm faster at every measured churn up to {pts[-1][0]:.0%} (ratio there {pts[-1][1]:.2f})"}
    if all(v <= 1 for _, v in pts):
        return {"churn": 0.0, "note": "CPU arm faster at every measured churn"}
    for (c0, v0), (c1, v1) in zip(pts, pts[1:]):
        if v0 > 1 >= v1:
            t = math.log(v0) / (math.log(v0) - math.log(v1))
            c = math.exp(math.log(c0) + t * (math.log(c1) - math.log(c0)))
            return {"churn": c, "note": f"between the measured {c0:.1%} and {c1:.1%}"}
    return {"churn": None, "note": "not monotonic in the measured range"}
