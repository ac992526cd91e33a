"""Arena snapshots (ncd_arena_save / ncd_arena_load) and the offline diff built on them:
a saved and re-loaded cluster must scan exactly like the live one (and like the oracle), and a
damaged file must be refused, not scanned."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle
from nexus_configuration_controller_b200 import arena as A
from nexus_configuration_controller_b200._ffi import NcdError
from nexus_configuration_controller_b200.drift import Arena
from nexus_configuration_controller_b200.offline_diff import offline_diff
from tests.helpers import random_clusters

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def save_all(ctx, tmp_path, ctrl, shards):
    paths = []
    for name, h in [("ctrl", ctrl)] + [(f"shard{s}", h) for s, h in enumerate(shards)]:
        a = ctx.arena_from_host(h)
        p = str(tmp_path / f"{name}.ncda")
        a.save(p)
        a.close()
        paths.append(p)
    return paths


@pytest.mark.parametrize("seed", [1, 2])
def test_saved_arenas_scan_like_the_live_ones(ctx, tmp_path, seed):
    ctrl, shards = random_clusters(seed + 60, n_templates=16, n_shards=3)
    want_codes, want_ents = oracle.scan(ctrl, shards)
    paths = save_all(ctx, tmp_path, ctrl, shards)
    for device_only in (True, False):
        ca = Arena.load(ctx, paths[0], device_only=device_only)
        sas = [Arena.load(ctx, p, device_only=device_only) for p in paths[1:]]
        for mode in (A.MODE_DIRECT, A.MODE_HASH64):
            d = ctx.scan(ca, sas, mode)
            np.testing.assert_array_equal(d.codes(), want_codes)
            np.testing.assert_array_equal(d.entries(), want_ents)
            d.close()
        # the loaded arena is the saved one, array for array
        back = ca.download()
        for name, _ in A.HostArena.FIELDS:
            np.testing.assert_array_equal(getattr(back, name), getattr(ctrl, name), err_msg=name)
        for a in sas + [ca]:
            a.close()
    # a HOST_PINNED load can take value syncs afterwards: its pinned arrays hold the snapshot
    ca = Arena.load(ctx, paths[0], device_only=False, payload_slack=1 << 16)
    arr = ca.host_arrays()
    i = next(k for k in range(ctrl.nobj) if ctrl.rec_len[k] > 0)
    arr["payload"][int(arr["rec_off"][i])] ^= 0xFF
    ca.sync_records(np.array([i], dtype=np.uint32), ctrl.payload_bytes)
    assert ca.download().record(i) != ctrl.record(i)
    ca.close()


def test_offline_diff_tool(ctx, tmp_path):
    ctrl, shards = random_clusters(77, n_templates=10, n_shards=2)
    _, want_ents = oracle.scan(ctrl, shards)
    paths = save_all(ctx, tmp_path, ctrl, shards)
    out = offline_diff(ctx, paths[0], paths[1:], orphans=True, names=True)
    assert [(e["object"], e["dependent"], e["shard"]) for e in out["entries"]] == [tuple(int(x) for x in r[:3]) for r in want_ents]
    assert all(e["key"] == ctrl.key(e["object"]) for e in out["entries"])
    ckeys = {(int(ctrl.kind[i]), ctrl.key(i)) for i in range(ctrl.nobj)}
    want_orphans = [{"shard": s, "object": i} for s, sh in enumerate(shards) for i in range(sh.nobj) if (int(sh.kind[i]), sh.key(i)) not in ckeys]
    assert out["orphans"] == want_orphans
    # the same through the command line, in a fresh process
    res = subprocess.run([sys.executable, "-m", "nexus_configuration_controller_b200.offline_diff", *paths, "--mode", "xxh64"],
                         cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr
    cli = json.loads(res.stdout)
    assert [(e["object"], e["dependent"], e["shard"], e["code"]) for e in cli["entries"]] == \
           [(e["object"], e["dependent"], e["shard"], e["code"]) for e in out["entries"]]


def test_damaged_snapshots_are_refused(ctx, tmp_path):
    ctrl, shards = random_clusters(5, n_templates=8, n_shards=1)
    path = save_all(ctx, tmp_path, ctrl, shards)[0]
    blob = open(path, "rb").read()
    cases = {
        "flipped payload byte": blob[:200] + bytes([blob[200] ^ 1]) + blob[201:],
        "truncated": blob[: len(blob) // 2],
        "no trailer": blob[:-16],
        "wrong magic": b"NOTARENA" + blob[8:],
        "empty": b"",
    }
    for what, data in cases.items():
        p = str(tmp_path / "bad.ncda")
        open(p, "wb").write(data)
        with pytest.raises(NcdError):
            Arena.load(ctx, p)
    with pytest.raises(NcdError):
        Arena.load(ctx, str(tmp_path / "missing.ncda"))
    # a file whose checksum is right but whose contents are not a valid arena is caught by the
    # same validation a commit runs: record 0 moved outside the payload, checksum recomputed
    import struct

    M = (1 << 64) - 1

    def mix(h, data):
        n = len(data) // 8 * 8
        for (w,) in struct.iter_unpack("<Q", data[:n]):
            h = ((h ^ w) * 0x9E3779B185EBCA87) & M
            h ^= h >> 29
        if n < len(data):
            w = int.from_bytes(data[n:], "little")
            h = ((h ^ w ^ ((len(data) - n) << 56)) * 0x9E3779B185EBCA87) & M
            h ^= h >> 29
        return h

    hdr = 8 + 4 + 4 + 5 * 8
    pos, h, fields = hdr, 0x6E6364615F736E61, []
    for _ in range(17):
        (nbytes,) = struct.unpack_from("<Q", blob, pos)
        fields.append((pos + 8, nbytes))
        h = mix(h, blob[pos + 8 : pos + 8 + nbytes])
        pos += 8 + (nbytes + 15) // 16 * 16
    assert struct.unpack_from("<Q", blob, pos)[0] == h  # the Python restatement of the checksum agrees
    bad = bytearray(blob)
    off1 = fields[1][0]  # rec_off array
    struct.pack_into("<Q", bad, off1, 1 << 40)
    h = 0x6E6364615F736E61
    for start, nbytes in fields:
        h = mix(h, bytes(bad[start : start + nbytes]))
    struct.pack_into("<Q", bad, pos, h)
    p = str(tmp_path / "bad.ncda")
    open(p, "wb").write(bytes(bad))
    with pytest.raises(NcdError):
        Arena.load(ctx, p)
    a = Arena.load(ctx, path)  # and the untouched file still loads
    assert a.download().nobj == ctrl.nobj
    a.close()
