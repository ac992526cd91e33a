"""What the synthetic generator was asked to inject, restated in plain Python.

Used to check a scan of synthetic arenas WITHOUT reading their payload: the decision for
(group g, member j, shard s) follows from the drift function alone, whatever the sizes — the
size-independent property behind the full-size tests.  Mirrors the rules of
ncd_synth.cu (drift_of, make_key) and ncd_keyhash of include/ncdrift.h.
"""
from __future__ import annotations

import numpy as np

M64 = (1 << 64) - 1
PPM = 1_000_000
SKIP, CREATE, UPDATE_DATA, UPDATE_OWNER, ERR_ROGUE = 0, 1, 2, 4, 8


def splitmix64(x: int) -> int:
    x = (x + 0x9E3779B97F4A7C15) & M64
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & M64
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & M64
    return x ^ (x >> 31)


def fmix64(x: int) -> int:
    x ^= x >> 33
    x = (x * 0xFF51AFD7ED558CCD) & M64
    x ^= x >> 33
    x = (x * 0xC4CEB9FE1A85EC53) & M64
    x ^= x >> 33
    return x


def keyhash(kind: int, key: bytes) -> int:
    h = 0xCBF29CE484222325
    h = ((h ^ (kind & 0xFF)) * 0x100000001B3) & M64
    for b in key:
        h = ((h ^ b) * 0x100000001B3) & M64
    return fmix64(h)


def template_key(g: int) -> bytes:
    return b"nexus/t-%010d" % g


def kept_groups(cfg) -> list:
    if cfg.part_mod <= 1:
        return list(range(cfg.groups))
    return [g for g in range(cfg.groups) if keyhash(0, template_key(g)) % cfg.part_mod == cfg.part_rem]


def drift(cfg, gid: int, j: int, s: int):
    d = splitmix64((cfg.drift_seed + gid * 0x9E3779B97F4A7C15 + s * 0xD1B54A32D192ED03) & M64)
    u = d % PPM
    missing = u < cfg.p_missing_ppm
    flip = (not missing) and u < cfg.p_missing_ppm + cfg.p_flip_ppm
    longer = (not missing) and (not flip) and u < cfg.p_missing_ppm + cfg.p_flip_ppm + cfg.p_len_ppm
    rogue = foreign = False
    if j > 0:
        u2 = splitmix64(d ^ 0x5851F42D4C957F2D) % PPM
        rogue = u2 < cfg.p_rogue_ppm
        foreign = (not rogue) and u2 < cfg.p_rogue_ppm + cfg.p_owner_ppm
    return missing, flip or longer, rogue, foreign


def expected_codes(cfg, nshards: int) -> np.ndarray:
    """uint8[rows, nshards] for the groups this partition keeps (rows = groups * (1 + deps))."""
    per = 1 + cfg.n_secrets + cfg.n_configmaps
    groups = kept_groups(cfg)
    out = np.zeros((len(groups) * per, nshards), dtype=np.uint8)
    for k, g in enumerate(groups):
        for s in range(nshards):
            t_missing, t_diff, _, _ = drift(cfg, g * per, 0, s)
            out[k * per, s] = CREATE if t_missing else (UPDATE_DATA if t_diff else SKIP)
            for j in range(1, per):
                missing, diff, rogue, foreign = drift(cfg, g * per + j, j, s)
                if missing:
                    code = CREATE
                elif rogue:
                    code = ERR_ROGUE
                else:
                    # a template that has to be created on the shard cannot be referenced yet
                    code = (UPDATE_DATA if diff else 0) | (UPDATE_OWNER if (foreign or t_missing) else 0)
                out[k * per + j, s] = code
    return out


def expected_diff_count(cfg, nshards: int) -> int:
    return int(np.count_nonzero(expected_codes(cfg, nshards)))


# ---------------------------------------------------------------- vectorised (full-size configs)
def _sm64(x: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return x ^ (x >> np.uint64(31))


def expected_codes_np(cfg, nshards: int) -> np.ndarray:
    """Same as expected_codes for an UNPARTITIONED config, in numpy (millions of rows)."""
    assert cfg.part_mod <= 1
    per = 1 + cfg.n_secrets + cfg.n_configmaps
    n = cfg.groups * per
    gid = np.arange(n, dtype=np.uint64)
    j = (gid % np.uint64(per)).astype(np.int64)
    out = np.zeros((n, nshards), dtype=np.uint8)
    pm, pf, pl = cfg.p_missing_ppm, cfg.p_flip_ppm, cfg.p_len_ppm
    for s in range(nshards):
        with np.errstate(over="ignore"):
            d = _sm64(np.uint64(cfg.drift_seed) + gid * np.uint64(0x9E3779B97F4A7C15) + np.uint64((s * 0xD1B54A32D192ED03) & M64))
        u = d % np.uint64(PPM)
        missing = u < pm
        diff = (~missing) & (u < pm + pf + pl)
        u2 = _sm64(d ^ np.uint64(0x5851F42D4C957F2D)) % np.uint64(PPM)
        dep = j > 0
        rogue = dep & (u2 < cfg.p_rogue_ppm)
        foreign = dep & ~rogue & (u2 < cfg.p_rogue_ppm + cfg.p_owner_ppm)
        t_missing = np.repeat(missing[::per], per)  # the group's template is absent on this shard
        code = np.where(diff, UPDATE_DATA, 0).astype(np.uint8)
        code = np.where(dep & (foreign | t_missing), code | UPDATE_OWNER, code)
        code = np.where(rogue, ERR_ROGUE, code)
        code = np.where(missing, CREATE, code)
        out[:, s] = code
    return out
