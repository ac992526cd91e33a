"""The C++ arena oracle against a dict-and-bytes restatement of the handlers in plain Python (CPU)."""
import numpy as np

import oracle
from nexus_configuration_controller_b200 import arena as A
from tests.helpers import random_clusters


def py_scan(ctrl: A.HostArena, shards):
    """controller.go:504-626, 720-745, 790-831 with dicts for the listers and == for DeepEqual."""

    def store(a):
        s = {}
        for i in range(a.nobj):
            s.setdefault((int(a.kind[i]), a.key(i)), i)
        return s

    def aux(a, off, ln):
        return a.aux[int(off) : int(off) + int(ln)].tobytes()

    cstore, sstores = store(ctrl), [store(s) for s in shards]
    rows = A.rows_of(ctrl)
    codes = np.zeros((len(rows), len(shards)), dtype=np.uint8)
    for r, (t, dep) in enumerate(rows):
        for s, sh in enumerate(shards):
            st = sstores[s].get((int(ctrl.kind[t]), ctrl.key(t)))
            if dep == 0:
                if st is None:
                    codes[r, s] = A.CREATE
                else:
                    codes[r, s] = A.SKIP if ctrl.record(t) == sh.record(st) else A.UPDATE_DATA
                continue
            d = int(ctrl.dep_start[t]) + dep - 1
            dkey = (int(ctrl.dep_kind[d]), aux(ctrl, ctrl.dep_key_off[d], ctrl.dep_key_len[d]).decode())
            cd = cstore.get(dkey)
            if cd is None:
                codes[r, s] = A.ERR_MISSING_REF
                continue
            sd = sstores[s].get(dkey)
            if sd is None:
                codes[r, s] = A.CREATE
                continue
            o0, o1 = int(sh.own_start[sd]), int(sh.own_start[sd + 1])
            if o0 == o1:
                codes[r, s] = A.ERR_ROGUE
                continue
            missing = True
            if st is not None:
                uid = aux(sh, sh.uid_off[st], sh.uid_len[st])
                for o in range(o0, o1):
                    if sh.own_is_template[o] and aux(sh, sh.own_uid_off[o], sh.own_uid_len[o]) == uid:
                        missing = False
            data = 0 if ctrl.record(cd) == sh.record(sd) else A.UPDATE_DATA
            codes[r, s] = data | (A.UPDATE_OWNER if missing else 0)
    return codes


def test_cpp_oracle_matches_python_restatement():
    seen = set()
    for seed in range(1, 9):
        ctrl, shards = random_clusters(seed, n_templates=8 + seed, n_shards=1 + seed % 4, max_len=3000)
        codes, ents = oracle.scan(ctrl, shards)
        want = py_scan(ctrl, shards)
        np.testing.assert_array_equal(codes, want)
        rows = A.rows_of(ctrl)
        want_ents = [(rows[r][0], rows[r][1], s, want[r, s]) for r in range(len(rows)) for s in range(len(shards)) if want[r, s]]
        np.testing.assert_array_equal(ents, np.array(want_ents, dtype=np.uint32).reshape(-1, 4))
        seen |= set(np.unique(codes).tolist())
    assert seen == {0, 1, 2, 4, 6, 8, 16}


def test_threaded_oracle_is_identical():
    ctrl, shards = random_clusters(5, n_templates=80, n_shards=4, max_len=2000)
    c1, e1 = oracle.scan(ctrl, shards, nthreads=1)
    c4, e4 = oracle.scan(ctrl, shards, nthreads=4)
    np.testing.assert_array_equal(c1, c4)
    np.testing.assert_array_equal(e1, e4)
    sec, c, nbytes = oracle.scan_timed(ctrl, shards, nthreads=2, reps=2)
    np.testing.assert_array_equal(c1, c)
    assert sec > 0 and nbytes > 0
