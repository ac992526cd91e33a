#!/usr/bin/env python
"""bench.py — the drift-check hot path on N B200s, one JSON line on stdout.

Metric (BASELINE.json): objects diffed per second — decisions (row x shard) per second — with
the bytes/s of the dominant kernel against the measured HBM roofline beside it.

Workload at N=1: BASELINE.json configs[3], the one the north_star quotes its target on and which
fits one GPU: 1M mixed objects (templates, Secrets, ConfigMaps, canonical length uniform in
1..16 KiB) x 8 shards, 0.1 % same-length byte drift + 0.01 % missing, ~78 GB resident in HBM.
At N>1 the same 1M objects are partitioned by ncd_keyhash(template key) % N (strong scaling,
as the north_star states it) and the per-GPU diff lists are exchanged with ONE NCCL all-gather.

A step = one full scan (join + compare + classify + compaction [+ exchange]) over inputs
already resident in HBM.  Inputs (>= 9.8 GB per GPU) are far larger than the 126 MB L2, so no
flush is needed between steps.

`e2e` is the same metric through the API the controller would call — the DriftChecker of
INTEGRATION.md (libncdhost.so): per step one interval's informer events (1 % of the controller's
objects changed, every shard catching up with the previous interval) are canonicalised and applied,
the changed records are packed into pinned memory and pulled over PCIe, every pair is scanned, the
diff list is read back, the sweep queue and a few lookups are served; host wall clock around the
whole call sequence, with the pack / upload / scan split and the bytes moved (bench_steady.py).
`e2e_by_churn` repeats it at 0.1 %, 1 %, 10 % and 100 % churn, `e2e_cold` is the stateless extreme
(every arena uploaded and indexed again every step through the C ABI), `break_even_churn` says where
the GPU arm stops beating the CPU arm.

--impl reference times the CPU arm: the reference's handlers restated over parsed objects (map
lookups + reflect.DeepEqual; the Go reference cannot be built in this image), fed the same events,
on the box's host cores, on a bounded sample.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "objects_diffed_per_s"
UNIT = "decisions/s"

WORKLOADS = {
    # name: (groups, n_secrets, n_configmaps, tmpl_lo, tmpl_hi, dep_lo, dep_hi, shards, flip_ppm, missing_ppm)
    "cfg2_10k_templates_4KB_x8_nodrift": (10_000, 1, 1, 512, 512, 4096, 4096, 8, 0, 0),
    "cfg3_100k_configmaps_64KB_x16_1pct": (100_000, 0, 1, 512, 512, 65536, 65536, 16, 10_000, 0),
    "cfg4_1M_mixed_1-16KB_x8": (333_334, 1, 1, 1024, 16384, 1024, 16384, 8, 1000, 100),
    "cfg5_10k_secrets_1MB_x4": (10_000, 1, 0, 512, 512, 1 << 20, 1 << 20, 4, 0, 0),
}
DEFAULT_WORKLOAD = "cfg4_1M_mixed_1-16KB_x8"


def make_cfg(name: str, groups: int | None = None, part_mod: int = 0, part_rem: int = 0):
    from nexus_configuration_controller_b200._ffi import SynthCfg

    g, ns, nc, tl, th, dl, dh, shards, flip, missing = WORKLOADS[name]
    c = SynthCfg()
    c.groups = groups if groups is not None else g
    c.n_secrets, c.n_configmaps = ns, nc
    c.tmpl_len_lo, c.tmpl_len_hi, c.dep_len_lo, c.dep_len_hi = tl, th, dl, dh
    cfg_id = list(WORKLOADS).index(name) + 2
    c.seed, c.drift_seed = 0xC0FFEE ^ cfg_id, 0xD21F7 ^ cfg_id
    c.p_flip_ppm, c.p_missing_ppm, c.p_len_ppm, c.p_owner_ppm, c.p_rogue_ppm = flip, missing, 0, 0, 0
    c.part_mod, c.part_rem, c.shuffle = part_mod, part_rem, 1
    # experiment knobs (never set by the driver's default invocation)
    if os.environ.get("NCD_BENCH_SHUFFLE"):
        c.shuffle = int(os.environ["NCD_BENCH_SHUFFLE"])
    if os.environ.get("NCD_BENCH_LEN"):
        lo, hi = [int(x) for x in os.environ["NCD_BENCH_LEN"].split(",")]
        c.tmpl_len_lo, c.tmpl_len_hi, c.dep_len_lo, c.dep_len_hi = lo, hi, lo, hi
    return c, shards


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """SM clock and throttle reasons sampled WHILE the timed region runs (NVML every ~2 ms; the
    timed region is only ~0.1 s, too short for an nvidia-smi process per sample)."""

    def __init__(self, gpu_index: int):
        self.gpu, self.samples, self.reason_bits, self.stop_flag, self.th = gpu_index, [], 0, threading.Event(), None
        self.max_mhz, self.nv, self.h = None, None, None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            # NVML indexes physical devices; honour CUDA_VISIBLE_DEVICES when it is a plain index list
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            phys = int(vis.split(",")[gpu_index]) if vis and all(x.strip().isdigit() for x in vis.split(",")) else gpu_index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        while not self.stop_flag.is_set():
            try:
                if nv is not None:
                    self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                    get = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
                    self.reason_bits |= int(get(self.h))
                else:
                    out = subprocess.run(
                        ["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm", "--format=csv,noheader,nounits", "-i", str(self.gpu)],
                        capture_output=True, text=True, timeout=5).stdout.strip().split(",")
                    self.samples.append(float(out[0]))
                    self.max_mhz = float(out[1])
            except Exception:
                pass
            self.stop_flag.wait(0.002)

    def start(self):
        self.th = threading.Thread(target=self._run, daemon=True)
        self.th.start()

    def stop(self) -> dict:
        self.stop_flag.set()
        if self.th:
            self.th.join(timeout=6)
        reasons = []
        if self.nv is not None:
            names = {
                "hw_slowdown": "nvmlClocksEventReasonHwSlowdown", "hw_thermal_slowdown": "nvmlClocksEventReasonHwThermalSlowdown",
                "sw_thermal_slowdown": "nvmlClocksEventReasonSwThermalSlowdown", "sw_power_cap": "nvmlClocksEventReasonSwPowerCap",
                "hw_power_brake_slowdown": "nvmlClocksEventReasonHwPowerBrakeSlowdown",
            }
            for label, const in names.items():
                bit = getattr(self.nv, const, None)
                if bit is None:
                    bit = getattr(self.nv, const.replace("ClocksEventReason", "ClocksThrottleReason"), 0)
                if bit and self.reason_bits & int(bit):
                    reasons.append(label)
        sm = sorted(self.samples)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_min_mhz": sm[0] if sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvml" if self.nv is not None else "nvidia-smi"}


def api_bound_model(templates: int, shards: int, writes: float) -> dict:
    """The reference's end-to-end sweep is bound by its own admission limiter and by REST round
    trips, not by the drift check: 50 items/s, burst 300 (controller.go:257-260, .helm/values.yaml
    :165-169), one HTTPS write per drifted object (assumed 5 ms in-region).  No Go toolchain or API
    server exists here, so this is a model, not a measurement."""
    rate, burst, rtt_s = 50.0, 300, 0.005
    admit_s = max(0.0, (templates - burst) / rate)
    return {
        "kind": "model (not measured: no Go toolchain, no API server on this box)",
        "limiter": "50 items/s, burst 300 (controller.go:257-260)",
        "sweep_admission_s": round(admit_s, 1),
        "rest_writes": int(writes), "rest_write_s_at_5ms_sequential": round(writes * rtt_s, 2),
        "note": f"a sweep of {templates} templates over {shards} shards takes >= {admit_s:.0f} s in the reference whatever the "
                "compare costs; the GPU path moves the compare, not this bound, unless the sweep is batched (DESIGN.md section 9)",
    }


def gathered_matches_spec(parts, workload, groups, S, world) -> bool:
    """N > 1: the gathered diff lists, mapped back from every rank's local object numbering to the object
    numbering of the UN-partitioned config, against the decisions the generator's drift function implies
    (tests/synth_spec.py: plain numpy, no arenas, no payload) — content, not only the count."""
    import numpy as np

    from tests import synth_spec as SS

    cfg, _ = make_cfg(workload, groups=groups)
    per = 1 + cfg.n_secrets + cfg.n_configmaps
    want = SS.expected_codes_np(cfg, S)
    # rank of every group: ncd_keyhash(TEMPLATE, "nexus/t-%010d") % world, vectorised FNV-1a + fmix64
    g = np.arange(cfg.groups, dtype=np.uint64)
    digits = [(g // np.uint64(10 ** (9 - k))) % np.uint64(10) + np.uint64(48) for k in range(10)]
    h = np.full(cfg.groups, 0xCBF29CE484222325, dtype=np.uint64)
    P = np.uint64(0x100000001B3)
    with np.errstate(over="ignore"):
        h = (h ^ np.uint64(0)) * P  # kind byte (template)
        for b in b"nexus/t-":
            h = (h ^ np.uint64(b)) * P
        for d in digits:
            h = (h ^ d) * P
        h ^= h >> np.uint64(33)
        h *= np.uint64(0xFF51AFD7ED558CCD)
        h ^= h >> np.uint64(33)
        h *= np.uint64(0xC4CEB9FE1A85EC53)
        h ^= h >> np.uint64(33)
    rank_of = (h % np.uint64(world)).astype(np.int64)
    got = np.zeros_like(want)
    for r, (ns, rows, ents) in enumerate(parts):
        if ns != S or len(ents) == 0:
            if ns != S:
                return False
            continue
        gsel = np.nonzero(rank_of == r)[0]
        if rows != len(gsel) * per or np.any(ents[:, 0] % per != 0):
            return False
        grow = gsel[ents[:, 0] // per] * per + ents[:, 1]
        got[grow, ents[:, 2]] = ents[:, 3].astype(np.uint8)
    return bool(np.array_equal(got, want))


def ncu_traffic(workload: str, kernel: str, n_gpus: int, default_size: bool):
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture, if one matches."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)[workload][kernel]
        if default_size and t["n_gpus"] == n_gpus:
            return t["dram_bytes_per_launch"], t["source"]
    except Exception:
        pass
    return None, None


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


# ----------------------------------------------------------------------------- CPU arm
def deep_equal_walk_baseline(mean_len: int, threads: int, budget_s: float = 2.0) -> dict:
    """Second CPU figure (SURVEY section 8d asks for both): reflect.DeepEqual restated on the value
    model, walking `map[string][]byte{"token": <mean_len bytes>}` pairs that are already parsed and
    separately allocated (a sweep streams from memory, as over an informer cache).  No lookup, no
    ownership logic: compare it with the memcmp figure as the cost of the walk alone."""
    import struct

    import oracle

    data = os.urandom(mean_len)
    wire = (b"\x08" + struct.pack("<I", 1) + b"\x03" + struct.pack("<I", 5) + b"token" +
            b"\x04" + struct.pack("<I", len(data)) + data)
    npairs = max(1024, min(1 << 16, (256 << 20) // max(2 * mean_len, 1)))  # ~256 MB of values: beyond the caches
    sec, eq = oracle.deep_equal_timed(wire, wire, npairs, 1, threads)  # warm-up, and sizes the repetitions
    reps = max(1, min(64, int(budget_s / max(sec, 1e-4))))
    sec, eq = oracle.deep_equal_timed(wire, wire, npairs, reps, threads)
    return {"value": npairs * reps / sec, "unit": "pairs/s", "cores": threads, "gb_per_s": 2 * mean_len * npairs * reps / sec / 1e9,
            "all_equal": bool(eq),
            "sample": f"{npairs} pairs of map[string][]byte with one {mean_len}-byte value, {reps} sweep(s); "
                      f"oracle DeepEqual on parsed values, no lookups"}


STEADY_LEVELS = [0.001, 0.01, 0.1, 1.0]
HEADLINE_CHURN = 0.01


def sized_groups(want: int, shards: int, copies: float) -> int:
    """Shrinks a sample so that `copies` host copies of it fit comfortably in the box's free memory."""
    import bench_steady as BS

    per_group = 3 * 8704 * (1 + shards)
    fit = int(0.6 * BS.mem_available_gb() * 1e9 / (per_group * copies))
    return max(1000, min(want, fit))


class CpuArm:
    """The reference's walk restated on the CPU (the Go reference cannot be built here), on a bounded sample of
    `workload`, fed the SAME informer events as the GPU arm's end-to-end leg (bench_steady.py): every cluster's
    cache as parsed objects, events parsed and stored, then a full sweep of every template and workgroup over
    every shard with map lookups and reflect.DeepEqual on the parsed values (oracle fleet).  The sample comes
    from the same generator as the GPU arm's input (on the GPU, then copied to host memory)."""

    def __init__(self, workload: str, sample_groups: int, threads: int, with_gpu_check: bool = True):
        import bench_steady as BS

        self.BS, self.workload, self.threads = BS, workload, threads
        _, S = make_cfg(workload)
        self.groups = sized_groups(sample_groups, S, 3.0)
        cfg, self.S = BS.converged_cfg(make_cfg, workload, self.groups)
        self.device = int(os.environ.get("LOCAL_RANK", "0"))
        self.hosts = BS.download_clusters(self.device, cfg, self.S)
        ctrl = self.hosts[0]
        self.algo = int(ctrl.rec_len[: ctrl.nobj].astype("uint64").sum()) * (1 + self.S)
        self.mean_len = max(1, self.algo // max(1, ctrl.nobj * (1 + self.S)))
        self.cpu = BS.SteadyCpu(self.hosts, threads)
        # the GPU arm on the same sample and the same events, outside any timed region: parity of the two arms
        self.gpu = BS.SteadyGpu(self.hosts, self.device, threads) if with_gpu_check else None
        self.rows = None

    def run(self, levels=None, steps_for=None) -> dict:
        BS = self.BS
        self.rows = BS.run_levels(self.gpu, self.cpu, self.hosts, levels or STEADY_LEVELS, steps_for or BS.default_steps, self.threads)
        return self.rows

    def report(self) -> dict:
        rows = self.rows
        head = rows[HEADLINE_CHURN]
        ctrl = self.hosts[0]
        out = {
            "value": head["cpu_decisions_per_s"], "unit": UNIT, "cores": self.threads, "kind": "port",
            "sample": f"{self.workload} shapes, converged fleet: first {self.groups} groups = {ctrl.nobj} objects x {self.S} shards "
                      f"({self.algo / 1e9:.2f} GB of values as parsed objects); a step = apply one interval's informer events "
                      f"({HEADLINE_CHURN:.1%} of the controller's objects changed + the shards catching up: {head['events_per_step']} events) "
                      f"then sweep every template / workgroup over every shard: map lookups + reflect.DeepEqual restated on parsed values; "
                      f"median of {head['steps']} steps of {head['cpu_ms_per_step']:.1f} ms",
            "ms_per_step": head["cpu_ms_per_step"],
            "gb_per_s": self.algo / (head["cpu_ms_per_step"] / 1e3) / 1e9,
            "by_churn": {f"{c:g}": {"decisions_per_s": r["cpu_decisions_per_s"], "ms_per_step": r["cpu_ms_per_step"],
                                    "events_per_step": r["events_per_step"]} for c, r in rows.items()},
        }
        if self.gpu is not None:
            out["identical_to_gpu"] = all(r.get("identical_to_cpu", False) for r in rows.values())
        out["deep_equal_walk"] = deep_equal_walk_baseline(int(self.mean_len), self.threads)
        return out

    def close(self):
        self.cpu.close()
        if self.gpu is not None:
            self.gpu.close()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    threads = os.cpu_count() or 1
    t0 = time.time()
    arm = CpuArm(args.workload, args.cpu_sample_groups, threads)
    # --steps / --warmup are honoured for the headline churn; the other levels keep their short defaults
    import bench_steady as BS

    arm.run(steps_for=lambda c: (args.steps, args.warmup) if c == HEADLINE_CHURN else BS.default_steps(c))
    base = arm.report()
    arm.close()
    value = base["value"]
    _, S = make_cfg(args.workload)
    line = {
        "impl": "reference",
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": base["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": args.workload, "shards": S,
                   "note": "CPU port of the reference's handlers on all host threads (the Go reference cannot be built here: "
                           "no Go toolchain, dependencies un-vendored); a step = one interval's informer events applied to the "
                           "caches + one resync sweep over the bounded sample, the same step the GPU arm's e2e leg makes"},
        "cpu_baseline": base,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_s": time.time() - t0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ----------------------------------------------------------------------------- GPU arm
def run_ours(args):
    import numpy as np
    import torch

    from nexus_configuration_controller_b200 import arena as A
    from nexus_configuration_controller_b200.drift import Context

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            sys.exit("bench.py --gpus N>1 must be launched with torch.distributed.run (one rank per GPU)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    # hash modes re-hash every copy in every step (digests are otherwise cached per committed arena)
    mode = {"direct": A.MODE_DIRECT, "hash64": A.MODE_HASH64 | A.MODE_REHASH, "sha256": A.MODE_SHA256 | A.MODE_REHASH,
            "xxh64": A.MODE_XXH64 | A.MODE_REHASH}[args.mode]

    ctx = Context(local)
    cfg, S = make_cfg(args.workload, groups=args.groups, part_mod=world if world > 1 else 0, part_rem=rank)
    t_setup = time.time()
    ca = ctx.synth_arena(cfg, -1)
    sas = [ctx.synth_arena(cfg, s) for s in range(S)]
    ctx.sync()
    setup_s = time.time() - t_setup

    from nexus_configuration_controller_b200 import multigpu as M

    ext = torch.cuda.ExternalStream(ctx.stream, device=local)
    # size of the diff-list buffer (the all-gather's send buffer): twice what a first scan finds,
    # the same on every rank
    cap_entries = args.gather_cap
    if cap_entries is None:
        d0 = ctx.scan(ca, sas, mode)
        n0 = len(d0.entries())
        d0.close()
        if dist:
            t = torch.tensor([n0], dtype=torch.int64, device=f"cuda:{local}")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            n0 = int(t.item())
        cap_entries = 1 << max(12, (2 * n0 + 1024 - 1).bit_length())
    sendbuf = torch.zeros(M.buffer_words(cap_entries), dtype=torch.int32, device=f"cuda:{local}")
    recvbuf = torch.zeros(world * sendbuf.numel(), dtype=torch.int32, device=f"cuda:{local}") if world > 1 else None

    xchg, xchg_note = None, None
    if world > 1 and args.exchange == "peer":
        # peer memory needs CUDA IPC and peer access between all GPUs of the box; if any rank cannot
        # map its peers, every rank takes the NCCL all-gather instead and the JSON line says so
        try:
            xchg = M.PeerExchange(ctx, world, rank, cap_entries)
            xchg.connect_group()
            ok = 1
        except Exception as e:  # noqa: BLE001 - reported below
            ok, xchg_note = 0, f"peer exchange unavailable on rank {rank}: {e}"
        if not ok:  # connect_group fails on every rank or on none
            if xchg is not None:
                xchg.close()
            xchg = None
            print(f"[bench rank {rank}] {xchg_note}; using the NCCL all-gather", file=sys.stderr)

    def step():
        """One scan, fully asynchronous: the compaction kernel writes the diff list into the exchange
        slot (ncd_scan_into) and, at N>1, the exchange of the per-GPU lists is queued behind it: one
        kernel of peer-memory stores (ncd_scan_exchange), or with --exchange nccl one all-gather."""
        if xchg is not None:
            xchg.scan(ca, sas, mode)
        else:
            M.scan_and_gather(ctx, ca, sas, mode, sendbuf, recvbuf, ext, cap_entries)

    # warm-up
    for _ in range(max(args.warmup, 3) if not args.allow_short_warmup else args.warmup):
        step()
    # the clock sampler (a thread, NVML) starts BEFORE the barrier: started after it, rank 0 entered its timed region
    # milliseconds late and every other rank's clock ran while it waited for rank 0's first lists
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    torch.cuda.synchronize()
    e0.record(ext)
    for _ in range(args.steps):
        step()
    e1.record(ext)
    torch.cuda.synchronize()
    elapsed_ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    # per-scan phase timings of the timed steps (CUDA events recorded by the library on its stream)
    cmp_ms, tot_ms, launches = [], [], 0
    phases = {"join_ms": 0.0, "hash_ms": 0.0, "compare_ms": 0.0, "classify_ms": 0.0}
    kept = min(args.steps, 60)  # the library keeps the phase events of the last 64 scans
    for back in range(kept):
        st = ctx.stats_back(back)
        cmp_ms.append(st.compare_ms if mode == A.MODE_DIRECT else st.hash_ms)
        tot_ms.append(st.total_ms)
        for k in phases:
            phases[k] += getattr(st, k) / kept
    launches = (st.launches + (1 if xchg is not None else 0)) * args.steps  # every step launches the same kernels
    gathered = xchg.fetch() if xchg is not None else None
    hdr = gathered[rank, :4] if xchg is not None else sendbuf[:4].cpu().numpy().view(np.uint32)
    if hdr[3]:
        raise RuntimeError(f"diff list of {hdr[0]} entries exceeds --gather-cap {cap_entries}")
    assert int(hdr[0]) == st.ndiff
    gathered_total, gathered_ok, rank_scan_ms = None, None, None
    if dist:
        parts = M.unpack_gathered(gathered if xchg is not None else recvbuf.cpu().numpy().view(np.uint32), world, cap_entries)
        gathered_total = sum(len(e) for _, _, e in parts)
        gathered_ok = gathered_matches_spec(parts, args.workload, args.groups, S, world) if rank == 0 else None
        per_rank = [torch.zeros(1, device=f"cuda:{local}") for _ in range(world)]
        dist.all_gather(per_rank, torch.tensor([sum(tot_ms) / len(tot_ms)], device=f"cuda:{local}"))
        rank_scan_ms = [round(float(x.item()), 4) for x in per_rank]
        t = torch.tensor([elapsed_ms], device=f"cuda:{local}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed_ms = float(t.item())
        tot = torch.tensor([float(st.pairs), float(st.algo_bytes), float(st.ndiff), float(launches)], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(tot)
        pairs_all, algo_all, ndiff_all, launches_all = [float(x) for x in tot.tolist()]
        dist.barrier()
    else:
        pairs_all, algo_all, ndiff_all, launches_all = float(st.pairs), float(st.algo_bytes), float(st.ndiff), float(launches)

    result = None
    if rank == 0:
        ms_per_step = elapsed_ms / args.steps
        value = pairs_all / (ms_per_step / 1e3)
        peak, peak_src = hbm_peak()
        kern_ms = sum(cmp_ms) / len(cmp_ms)
        achieved = st.algo_bytes / (kern_ms / 1e3) / 1e9  # this rank's kernel over this rank's bytes
        kernel_name = ({0: "k_compare", 1: "k_compare_grouped<1>", 2: "k_compare_grouped<2>", 4: "k_compare_grouped<4>", 8: "k_compare_tma", 16: "k_compare_small", 32: "k_scan_small"}.get(st.compare_variant, "k_compare")
                       if mode == A.MODE_DIRECT else {"hash64": "k_h64_tiles", "sha256": "k_sha256", "xxh64": "k_xxh64"}[args.mode])
        traffic, traffic_src = ncu_traffic(args.workload, kernel_name, world, args.groups is None and not os.environ.get("NCD_BENCH_LEN"))
        result = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {
                "workload": args.workload, "mode": args.mode, "objects": int(cfg.groups) * (1 + cfg.n_secrets + cfg.n_configmaps),
                "shards": S, "partition": f"ncd_keyhash(template key) % {world}" if world > 1 else "none",
                "algorithmic_bytes": algo_all, "l2": "inputs >> 126 MB L2, no flush needed", "setup_s": round(setup_s, 2),
                "diff_entries": ndiff_all, "diff_entries_after_allgather": gathered_total,
                "gathered_identical": gathered_ok if world > 1 else None,
                "rank_scan_ms": rank_scan_ms if world > 1 else None,
                "exchange": None if world == 1 else ("peer-memory stores over NVLink (k_exchange)" if xchg is not None else "nccl all_gather_into_tensor" + (f" ({xchg_note})" if xchg_note else "")),
            },
            "gb_per_s": algo_all / (ms_per_step / 1e3) / 1e9,
            "roofline": {
                "bound": "hbm", "kernel": kernel_name,
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src,
                "traffic": traffic, "traffic_source": traffic_src, "algorithmic_bytes_per_launch": int(st.algo_bytes), "kernel_ms": kern_ms, "scan_ms": sum(tot_ms) / len(tot_ms),
                "phases_ms": {k: round(v, 4) for k, v in phases.items()},
            },
            "gpu_launches": int(launches_all),
            # reported separately, as the north_star asks: what bounds a real reconcile sweep is not this path
            "api_bound_model": api_bound_model(int(cfg.groups), S, ndiff_all),
            "clocks": clocks,
        }
    # end-to-end: every rank, bounded sample; CPU baseline: rank 0 at N=1 only
    if xchg is not None:
        dist.barrier()  # nobody unmaps a buffer a peer may still be writing
        xchg.close()
    for a in sas + [ca]:
        a.close()
    if not args.skip_e2e:
        if world > 1:
            e2e_cap = 1 << 18
            e2e_send = torch.zeros(M.buffer_words(e2e_cap), dtype=torch.int32, device=f"cuda:{local}")
            e2e_recv = torch.zeros(world * e2e_send.numel(), dtype=torch.int32, device=f"cuda:{local}")
            gather = lambda d: M.gather_diffs(d, e2e_send, e2e_recv, ext, e2e_cap)  # noqa: E731
        else:
            gather = lambda d: None  # noqa: E731
        e2e_cold = run_e2e_cold(ctx, args, mode & 0xFF, world, rank, dist, ext, gather)
        if rank == 0:
            result["e2e_cold"] = e2e_cold
    ctx.close()
    if not args.skip_e2e:
        steady = run_steady(args, world, rank, local, dist)
        if rank == 0:
            head = steady["rows"][HEADLINE_CHURN]
            result["e2e"] = {
                "value": head["decisions_per_s"], "unit": UNIT, "h2d_bytes_per_step": head["h2d_bytes_per_step"],
                "d2h_bytes_per_step": head["d2h_bytes_per_step"], "ms_per_step": head["ms_per_step"], "churn": HEADLINE_CHURN,
                "events_per_step": head["events_per_step"], "value_bytes_per_step": head["value_bytes_per_step"],
                "split_ms": head["split_ms"], "diff_entries": head["diff_entries"],
                "api": "DriftChecker of INTEGRATION.md (libncdhost.so over libncdrift.so): ncdh_apply_events -> ncdh_scan -> "
                       "ncdh_sweep_queue + ncdh_lookup",
                "sample": steady["sample"],
            }
            result["e2e_by_churn"] = {f"{c:g}": r for c, r in steady["rows"].items()}
            result["e2e_first_scan"] = steady["first_scan"]
    if rank == 0 and world == 1 and not args.skip_cpu:
        arm = CpuArm(args.workload, args.cpu_sample_groups, os.cpu_count() or 1)
        arm.run()
        result["cpu_baseline"] = arm.report()
        arm.close()
        if "e2e_by_churn" in result:
            import bench_steady as BS

            cpu_by = result["cpu_baseline"]["by_churn"]
            # level by level against the CPU arm at the same churn (the CPU arm's own cost grows with churn too)
            ratios = {k: r["decisions_per_s"] / cpu_by[k]["decisions_per_s"] for k, r in result["e2e_by_churn"].items() if k in cpu_by}
            result["e2e_vs_cpu_by_churn"] = {k: round(v, 3) for k, v in ratios.items()}
            result["break_even_churn"] = BS.break_even(ratios)
    if rank == 0:
        print(json.dumps(result), flush=True)
    if dist:
        dist.destroy_process_group()
    return 0


def run_e2e_cold(ctx, args, mode, world, rank, dist, ext, gather):
    """The same metric through the C ABI with HOST (pinned) arenas and NO resident state: every step uploads
    every arena of this rank (H2D), rebuilds its index, scans and reads the diff list back — what a controller
    that kept nothing on the GPU would pay per sweep (100 % churn without even the canonicaliser).  Timed with
    CUDA events on the library's stream around the whole region (host gaps included), barrier on both sides,
    max over ranks."""
    import torch

    from nexus_configuration_controller_b200._ffi import check
    from nexus_configuration_controller_b200.drift import Arena

    cfg, S = make_cfg(args.workload, groups=args.e2e_groups, part_mod=world if world > 1 else 0, part_rem=rank)
    hosts = []
    for r in range(-1, S):
        a = ctx.synth_arena(cfg, r)
        c = a.counts()
        h = Arena(ctx, c.payload_bytes, c.nobj, c.aux_bytes, c.nowner, c.ndep, device_only=False)
        v = h.host_view()
        check(ctx.lib.ncd_arena_download(a.handle, C.byref(v)))  # fills the pinned host arrays
        hosts.append((h, c))
        a.close()
    h2d_cold = sum(c.payload_bytes + c.aux_bytes + c.nobj * 29 + 8 + c.nowner * 9 + c.ndep * 9 for _, c in hosts)
    arenas = [h for h, _ in hosts]

    def total(x):
        if not dist:
            return float(x)
        t = torch.tensor([float(x)], dtype=torch.float64, device=f"cuda:{ctx.device}")
        dist.all_reduce(t)
        return float(t.item())

    def cold_step():
        ctx.commit_many(arenas, [(c.nobj, c.payload_bytes, c.aux_bytes, c.nowner, c.ndep) for _, c in hosts])
        d = ctx.scan(arenas[0], arenas[1:], mode)
        gather(d)
        ents = d.entries()
        st = ctx.last_stats()
        d.close()
        return st, ents

    reps = max(1, args.e2e_steps)
    for _ in range(1):
        out = cold_step()
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    evs[0].record(ext)
    for i in range(reps):
        out = cold_step()
        evs[i + 1].record(ext)
    torch.cuda.synchronize()
    per_step = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(reps))
    cold_ms = per_step[len(per_step) // 2]
    if dist:
        t = torch.tensor([cold_ms], device=f"cuda:{ctx.device}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        cold_ms = float(t.item())
    st, ents = out
    pairs, algo = total(st.pairs), total(st.algo_bytes)
    result = {
        "value": pairs / (cold_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": int(total(h2d_cold)),
        "d2h_bytes_per_step": int(total(ents.nbytes + 64)), "ms_per_step": cold_ms, "gb_per_s": algo / (cold_ms / 1e3) / 1e9,
        "sample": f"{args.workload}: first {args.e2e_groups} groups = {int(total(hosts[0][1].nobj))} objects x {S} shards in pinned "
                  f"host arenas{' split over ' + str(world) + ' ranks by key hash' if world > 1 else ''}; "
                  f"step = H2D of every arena + index build + scan + diff list D2H; median of {reps} steps",
    }
    for h in arenas:
        h.close()
    return result


def run_steady(args, world, rank, local, dist):
    """The end-to-end legs: steady-state sweeps through the DriftChecker API at several churn levels (bench_steady.py)."""
    import torch

    import bench_steady as BS

    _, S0 = make_cfg(args.workload)
    threads = max(1, (os.cpu_count() or 1) // world)
    groups = sized_groups(args.e2e_groups, S0, 4.5 * world)
    cfg, S = BS.converged_cfg(make_cfg, args.workload, groups, part_mod=world if world > 1 else 0, part_rem=rank)
    hosts = BS.download_clusters(local, cfg, S)
    for h in hosts[1:]:
        h.payload = __import__("numpy").zeros(16, dtype="uint8")  # shard events take their values from the controller's records
    arm = BS.SteadyGpu(hosts, local, threads)
    if dist:  # the ranks sweep in lockstep: one tiny all-reduce per step stands for the exchange of the diff lists
        flag = torch.zeros(1, device=f"cuda:{local}")
        step0 = arm.step

        def step(batches):
            t0 = time.perf_counter()
            _, tm, n = step0(batches)
            dist.all_reduce(flag)
            torch.cuda.synchronize()
            return time.perf_counter() - t0, tm, n

        arm.step = step
        dist.barrier()
    levels = [c for c in STEADY_LEVELS if c <= args.max_churn]
    rows = BS.run_levels(arm, None, hosts, levels, BS.default_steps, threads)
    decisions, nobj = float(arm.decisions), float(hosts[0].nobj)
    first = dict(arm.first, load_s=round(arm.load_s, 2))
    arm.close()
    if dist:
        t = torch.tensor([decisions, nobj], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(t)
        decisions, nobj = [float(x) for x in t.tolist()]
        for c, r in rows.items():  # max over ranks of the step time, sums of the bytes
            t = torch.tensor([r["ms_per_step"]], dtype=torch.float64, device=f"cuda:{local}")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            r["ms_per_step"] = float(t.item())
            t = torch.tensor([float(r[k]) for k in ("h2d_bytes_per_step", "d2h_bytes_per_step", "events_per_step", "value_bytes_per_step", "diff_entries", "records_synced")],
                             dtype=torch.float64, device=f"cuda:{local}")
            dist.all_reduce(t)
            for k, v in zip(("h2d_bytes_per_step", "d2h_bytes_per_step", "events_per_step", "value_bytes_per_step", "diff_entries", "records_synced"), t.tolist()):
                r[k] = int(v)
            r["decisions_per_s"] = decisions / (r["ms_per_step"] / 1e3)
    sample = (f"{args.workload} shapes, converged fleet: first {groups} groups = {int(nobj)} objects x {S} shards resident behind the "
              f"DriftChecker{' split over ' + str(world) + ' ranks by key hash' if world > 1 else ''}; a step = one interval's informer "
              f"events (a fraction `churn` of the controller's objects changed + every shard catching up with the previous interval) "
              f"applied through ncdh_apply_events (values canonicalised on {threads} threads), then ncdh_scan (changed records packed into "
              f"pinned memory, pulled over PCIe, full scan of every pair, diff list to the host), ncdh_sweep_queue and 64 ncdh_lookup; "
              f"host wall clock, median over the steps")
    return {"rows": rows, "sample": sample, "first_scan": first}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=list(WORKLOADS))
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="N>1: how the per-GPU diff lists reach every rank")
    ap.add_argument("--mode", default="direct", choices=["direct", "hash64", "sha256", "xxh64"])
    ap.add_argument("--groups", type=int, default=None, help="override the number of template groups (smaller runs)")
    ap.add_argument("--gather-cap", type=int, default=None, help="entries per rank in the all-gather buffer (default: sized from a first scan)")
    ap.add_argument("--cpu-sample-groups", type=int, default=30000, help="template groups of the CPU arm's sample")
    ap.add_argument("--max-churn", type=float, default=1.0, help="largest churn level of the end-to-end legs")
    ap.add_argument("--e2e-groups", type=int, default=100000)
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--skip-e2e", action="store_true")
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--allow-short-warmup", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
