"""An independent THIRD statement of the row-order contract, for the parts no reference test pins (table row
order after mixed migrations, archetype creation order, the `{}` table).  A list-based model written directly from the
reference source — Archetype::AllocateRow (component/archetype.hpp:358-367: push_back), Archetype::Remove (:408-442: the LAST
row takes the hole, pop_back), ComponentManager::InitEntity (manager.hpp:1059-1072: new entities go to `{}`),
GetOrCreateArchetype (:1852-1877: creation order) and MigrateEntity (:1893-1952: allocate in dst FIRST, then swap-and-pop
src) — sharing no code with oracle/.  It carries entity ids only; the oracle must agree with it on every sequence."""
import numpy as np
import pytest

from heliosengine_b200 import capi, worlds

SHORTS = ["Position", "Velocity", "Health", "Team", "Damping"]


class ListWorld:
    def __init__(self):
        self.tables, self.where = {}, {}  # frozenset of types -> rows (dicts keep CREATION order); entity -> its frozenset

    def rows(self, key):
        return self.tables.setdefault(key, [])  # GetOrCreateArchetype: appended to the archetype list when first needed

    def create(self, e):  # InitEntity: a row at the tail of `{}`
        self.rows(frozenset()).append(e)
        self.where[e] = frozenset()

    def _pop(self, e):  # Archetype::Remove: swap-and-pop
        rows = self.tables[self.where[e]]
        rows[rows.index(e)] = rows[-1]
        rows.pop()

    def migrate(self, e, key):  # MigrateEntity: no-op on the same archetype; AllocateRow in dst, THEN Remove from src
        if e not in self.where or key == self.where[e]:
            return
        self.rows(key).append(e)
        self._pop(e)
        self.where[e] = key

    def apply(self, kind, e, types):
        if e not in self.where:  # dead / stale id: Try* skip silently (the sequences use Try* for those)
            return
        have = self.where[e]
        if kind in (capi.CMD_ADD, capi.CMD_TRY_ADD):
            self.migrate(e, have | types)
        elif kind in (capi.CMD_REMOVE, capi.CMD_TRY_REMOVE):
            self.migrate(e, have - types)
        elif kind == capi.CMD_CLEAR:
            self.migrate(e, frozenset())
        else:  # Destroy / TryDestroy
            self._pop(e)
            del self.where[e]

    def state(self):
        return [(sorted(k), list(v)) for k, v in self.tables.items()]


def oracle_state(w):
    return [(sorted(a["types"]), [int(x) for x in a["entities"]]) for a in w.snapshot(columns=False)["archetypes"]]


@pytest.mark.parametrize("seed", range(8))
def test_oracle_agrees_with_the_list_model(oracle, seed):
    run_against_model(oracle, seed)


@pytest.mark.gpu
@pytest.mark.parametrize("seed", range(8))
def test_gpu_agrees_with_the_list_model(gpu, seed):
    """The CUDA path against the list model directly (not through the oracle)."""
    run_against_model(gpu, seed)


def run_against_model(backend, seed):
    rng = np.random.default_rng(seed)
    w, m = capi.World(backend), ListWorld()
    worlds.register_all(w, SHORTS)
    h = {s: capi.fnv1a(worlds.cname(s)) for s in SHORTS}
    for e in w.create_entities(int(rng.integers(3, 40))):  # CreateEntity x n
        m.create(int(e))
    stale = []
    for frame in range(6):
        live = list(m.where)
        spawned = [int(e) for e in w.reserve_entities(int(rng.integers(0, 5)))]
        n_cmd = int(rng.integers(1, 120))
        pool = live + spawned + stale[-6:]
        cmds = np.zeros(n_cmd, dtype=capi.COMMAND_DTYPE)
        payload = bytearray()
        specs = []
        for i in range(n_cmd):
            e = int(pool[rng.integers(0, len(pool))])
            kind = int(rng.choice([capi.CMD_TRY_ADD, capi.CMD_TRY_ADD, capi.CMD_TRY_REMOVE, capi.CMD_TRY_DESTROY, capi.CMD_TRY_REMOVE]))
            if kind != capi.CMD_TRY_DESTROY and rng.random() < 0.3:  # non-Try forms only on ids that cannot be dead yet
                kind = {capi.CMD_TRY_ADD: capi.CMD_ADD, capi.CMD_TRY_REMOVE: capi.CMD_TRY_REMOVE}[kind]
            sh = [str(x) for x in rng.choice(SHORTS, size=int(rng.integers(1, 4)), replace=False)] if kind != capi.CMD_TRY_DESTROY else []
            if kind == capi.CMD_ADD and (e in stale or any(s[1] == e and s[0] == capi.CMD_TRY_DESTROY for s in specs)):
                kind = capi.CMD_TRY_ADD  # the non-Try form asserts on a dead entity in the reference
            specs.append((kind, e, frozenset(h[s] for s in sh)))
            cmds[i]["kind"], cmds[i]["entity"] = kind, e
            cmds[i]["component_mask"] = w.mask(worlds.cname(s) for s in sh)
            cmds[i]["payload_offset"] = len(payload)
            if kind in (capi.CMD_ADD, capi.CMD_TRY_ADD):
                payload += bytes(sum(worlds.COMPONENTS[s][0] for s in sh))
            if kind == capi.CMD_TRY_DESTROY:
                stale.append(e)
        if rng.random() < 0.3:  # Clear through a live target
            k = int(rng.integers(0, n_cmd))
            if specs[k][1] not in stale:
                specs[k] = (capi.CMD_CLEAR, specs[k][1], frozenset())
                cmds[k]["kind"], cmds[k]["component_mask"] = capi.CMD_CLEAR, 0
        w.enqueue(cmds, bytes(payload))
        w.flush()
        for e in spawned:  # World::Flush: reserved ids first, in index order (world.hpp:957-963), then the queue, FIFO
            m.create(e)
        for kind, e, types in specs:
            m.apply(kind, e, types)
        assert oracle_state(w) == m.state(), (seed, frame)
    w.close()


def test_tail_and_refill_by_hand(oracle):
    """The smallest sequence that tells 'current last row fills the hole' from every other rule."""
    w, m = capi.World(oracle), ListWorld()
    worlds.register_all(w, SHORTS)
    P, H = worlds.cname("Position"), worlds.cname("Health")
    e = [int(x) for x in w.spawn_batch([P], 5, [np.zeros(10, np.float32)])]
    # destroy e1 -> e4 fills row 1; add Health to e4 -> leaves row 1, e3 (now last) fills it; destroy e0 -> e2 fills row 0
    cmds = np.zeros(3, dtype=capi.COMMAND_DTYPE)
    cmds["kind"] = [capi.CMD_DESTROY, capi.CMD_ADD, capi.CMD_DESTROY]
    cmds["entity"] = [e[1], e[4], e[0]]
    cmds["component_mask"] = [0, w.mask([H]), 0]
    w.enqueue(cmds, bytes(4))
    w.flush()
    got = oracle_state(w)
    assert got == [([], []), ([capi.fnv1a(P)], [e[2], e[3]]), (sorted([capi.fnv1a(P), capi.fnv1a(H)]), [e[4]])]
