"""Times the paths next to the scan on one B200 (rows f-2 and f-4 of SURVEY.md section 8):
commit of a resident arena (index build), value sync, orphan search.

    python profiles/measure_extras.py            # cfg 4 shapes, 333 334 groups (1M objects)

Prints one JSON line.  Host wall-clock around synchronous C-ABI calls (each call ends with a stream
synchronise), after one warm-up call of each kind.
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

import bench  # noqa: E402
from nexus_configuration_controller_b200.drift import Arena, Context  # noqa: E402
from nexus_configuration_controller_b200._ffi import check  # noqa: E402
import ctypes as C  # noqa: E402


def timed(fn, reps=3):
    fn()
    t = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        t.append((time.perf_counter() - t0) * 1e3)
    return sorted(t)[len(t) // 2]


def main():
    groups = int(sys.argv[1]) if len(sys.argv) > 1 else None
    cfg, S = bench.make_cfg(bench.DEFAULT_WORKLOAD, groups=groups)
    out = {"workload": bench.DEFAULT_WORKLOAD, "groups": int(cfg.groups), "shards": S}
    with Context(0) as ctx:
        ca = ctx.synth_arena(cfg, -1)
        sas = [ctx.synth_arena(cfg, s) for s in range(S)]
        ctx.sync()
        n_shard_objs = sum(a.counts().nobj for a in sas)
        out["orphans_ms"] = timed(lambda: ctx.find_orphans(ca, sas, cap=1 << 16))
        out["orphans_found"] = int(len(ctx.find_orphans(ca, sas, cap=1 << 16)))
        out["orphans_shard_objects"] = int(n_shard_objs)
        # one shard as a pinned, resident arena: full commit, structural re-commit, value sync
        c = sas[0].counts()
        h = Arena(ctx, c.payload_bytes + (1 << 20), c.nobj, c.aux_bytes, c.nowner, c.ndep, device_only=False)
        v = h.host_view()
        check(ctx.lib.ncd_arena_download(sas[0].handle, C.byref(v)))
        out["arena_objects"] = int(c.nobj)
        out["arena_payload_gb"] = c.payload_bytes / 1e9
        out["commit_full_ms"] = timed(lambda: h.commit(c.nobj, c.payload_bytes, c.aux_bytes, c.nowner, c.ndep), reps=2)
        out["recommit_structural_ms"] = timed(
            lambda: h.recommit(c.nobj, c.payload_bytes, c.aux_bytes, c.nowner, c.ndep, c.payload_bytes))
        arr = h.host_arrays()
        rng = np.random.default_rng(1)
        for frac in (0.001, 0.01):
            k = max(1, int(c.nobj * frac))
            idx = rng.choice(c.nobj, size=k, replace=False).astype(np.uint32)
            nbytes = int(((arr["rec_len"][idx].astype(np.int64) + 15) // 16 * 16).sum())
            ms = timed(lambda: h.sync_records(idx, c.payload_bytes))
            out[f"sync_{frac}_ms"] = ms
            out[f"sync_{frac}_records"] = k
            out[f"sync_{frac}_gb_per_s"] = nbytes / ms / 1e6
        h.close()
        for a in sas + [ca]:
            a.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
