"""Offline drift check between arena snapshots.

    python -m nexus_configuration_controller_b200.offline_diff controller.ncda shard-a.ncda shard-b.ncda
        [--mode direct|hash64|xxh64|sha256] [--device 0] [--orphans] [--names]

Loads every snapshot onto the GPU (ncd_arena_load: checksum, validation, index), runs the same
scan a live controller would (ncd_scan) and prints one JSON document: the diff entries
(controller object, dependent, shard, decision code) and, with --orphans, the shard objects the
controller snapshot does not hold.  Snapshots are written by ``Arena.save`` / ``ncd_arena_save``.
There is no CPU path: without the CUDA library and a GPU this tool fails.
"""
from __future__ import annotations

import argparse
import json
import sys

from . import arena as A
from .drift import Arena, Context

CODE_NAMES = {A.CREATE: "CREATE", A.UPDATE_DATA: "UPDATE_DATA", A.UPDATE_OWNER: "UPDATE_OWNER",
              A.UPDATE_DATA | A.UPDATE_OWNER: "UPDATE_DATA|UPDATE_OWNER", A.ERR_ROGUE: "ERR_ROGUE",
              A.ERR_MISSING_REF: "ERR_MISSING_REF"}
MODES = {"direct": A.MODE_DIRECT, "hash64": A.MODE_HASH64, "xxh64": A.MODE_XXH64, "sha256": A.MODE_SHA256}


def offline_diff(ctx: Context, ctrl_path: str, shard_paths, mode: int = A.MODE_DIRECT, orphans: bool = False,
                 names: bool = False) -> dict:
    ctrl = Arena.load(ctx, ctrl_path)
    shards = [Arena.load(ctx, p) for p in shard_paths]
    try:
        d = ctx.scan(ctrl, shards, mode)
        ents = d.entries()
        st = ctx.last_stats()
        d.close()
        host = ctrl.download() if names else None
        out = {
            "controller": ctrl_path, "shards": list(shard_paths), "rows": int(st.rows), "decisions": int(st.pairs),
            "compared_bytes": int(st.algo_bytes), "scan_ms": float(st.total_ms),
            "entries": [
                dict({"object": int(o), "dependent": int(dep), "shard": int(s), "code": CODE_NAMES.get(int(c), int(c))},
                     **({"key": host.key(int(o))} if names else {}))
                for o, dep, s, c in ents
            ],
        }
        if orphans:
            out["orphans"] = [{"shard": int(s), "object": int(i)} for s, i in ctx.find_orphans(ctrl, shards)]
        return out
    finally:
        for a in shards + [ctrl]:
            a.close()


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("controller")
    ap.add_argument("shards", nargs="*")
    ap.add_argument("--mode", default="direct", choices=sorted(MODES))
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--orphans", action="store_true")
    ap.add_argument("--names", action="store_true", help="add the controller object's namespace/name to every entry")
    args = ap.parse_args(argv)
    with Context(args.device) as ctx:
        json.dump(offline_diff(ctx, args.controller, args.shards, MODES[args.mode], args.orphans, args.names), sys.stdout)
        sys.stdout.write("\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())
