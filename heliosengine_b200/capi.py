"""ctypes bindings over the C ABI declared in include/helios_gpu.h.

`load_gpu()` binds heliosengine_b200/libhelios_gpu.so (hand-written sm_100a kernels).  There is no CPU
fallback: if the library is missing or no CUDA device is usable the calls raise.  The same wrapper classes
can be pointed at the test oracle (oracle/liboracle.so exports the same entry points with the prefix
``helios_oracle_``) — that is done only by tests/, bench.py's cpu_baseline legs and smoke().
"""
from __future__ import annotations

import ctypes as C
import os
import struct
from typing import Iterable, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(_HERE)

MAX_COMPONENTS = 64
MAX_QUERY_COMPONENTS = 8
TILE_ROWS = 1024
RESOURCE_BYTES = 4096

OK = 0
ERR_INVALID_ARGUMENT, ERR_CUDA, ERR_NOT_FOUND, ERR_CAPACITY, ERR_STATE, ERR_CYCLE = -1, -2, -3, -4, -5, -6

CMD_DESTROY, CMD_TRY_DESTROY, CMD_ADD, CMD_TRY_ADD, CMD_REMOVE, CMD_TRY_REMOVE, CMD_CLEAR = range(7)
EXEC_SERIAL, EXEC_STREAMS, EXEC_GRAPH = 0, 1, 2
MSG_AUTOMATIC, MSG_MANUAL = 0, 1
SYSTEM_EMITS_COMMANDS = 1
SYSTEM_NEEDS_VIEW = 2
SYSTEM_CONSUMES_MESSAGES = 4

WORLD_NO_BUILTIN_MESSAGES, WORLD_GENERAL_APPLY = 1, 2  # helios_gpu_world_config::flags
BUILTIN_ENTITY_ADDED, BUILTIN_ENTITY_DESTROYED, BUILTIN_COMPONENTS_CLEARED, BUILTIN_COMPONENT_ADDED, BUILTIN_COMPONENT_REMOVED = range(5)

INVALID_ENTITY = 0xFFFFFFFFFFFFFFFF

_FNV_BASIS = 14695981039346656037
_FNV_PRIME = 1099511628211


def fnv1a(name: str) -> int:
    """TypeIndex of a type name: FNV-1a-64 (reference utils/hash.hpp:26-32, utils/type_info.hpp:308-311)."""
    h = _FNV_BASIS
    for ch in name.encode():
        h = ((h ^ ch) * _FNV_PRIME) & 0xFFFFFFFFFFFFFFFF
    return h


def make_entity(index, generation=1):
    return (np.asarray(index, dtype=np.uint64) | (np.asarray(generation, dtype=np.uint64) << np.uint64(32)))


class WorldConfig(C.Structure):
    _fields_ = [("device", C.c_int32), ("entity_capacity", C.c_uint32), ("table_capacity", C.c_uint32), ("flags", C.c_uint32)]


class QueryDesc(C.Structure):
    _fields_ = [
        ("n_components", C.c_uint32),
        ("components", C.c_uint64 * MAX_QUERY_COMPONENTS),
        ("writes", C.c_uint8 * MAX_QUERY_COMPONENTS),
        ("n_with", C.c_uint32),
        ("with_", C.c_uint64 * MAX_COMPONENTS),
        ("n_without", C.c_uint32),
        ("without", C.c_uint64 * MAX_COMPONENTS),
        ("n_sparse_data", C.c_uint32),
        ("sparse_data", C.c_uint64 * MAX_QUERY_COMPONENTS),
        ("sparse_rule", C.c_uint32),
        ("n_optional", C.c_uint32),
        ("optional", C.c_uint64 * MAX_QUERY_COMPONENTS),
        ("optional_writes", C.c_uint8 * MAX_QUERY_COMPONENTS),
    ]


# helios_gpu_query_desc::sparse_rule (include/helios_gpu.h)
SPARSE_STRICT, SPARSE_FOREACH, SPARSE_WITH_ENTITY = 0, 1, 2


class Command(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("payload_offset", C.c_uint32), ("entity", C.c_uint64), ("component_mask", C.c_uint64)]


COMMAND_DTYPE = np.dtype([("kind", "<u4"), ("payload_offset", "<u4"), ("entity", "<u8"), ("component_mask", "<u8")])
assert COMMAND_DTYPE.itemsize == C.sizeof(Command) == 24
ROUTED_CMD_PAYLOAD = 40  # HELIOS_GPU_ROUTED_CMD_PAYLOAD
ROUTED_COMMAND_DTYPE = np.dtype([("kind", "<u4"), ("payload_bytes", "<u4"), ("entity", "<u8"), ("component_mask", "<u8"),
                                 ("payload", "u1", (ROUTED_CMD_PAYLOAD,))])
assert ROUTED_COMMAND_DTYPE.itemsize == 64
SPAWN_PLACEHOLDER_GENERATION = 0xFFFFFFFE


def spawn_placeholder(j):
    """HELIOS_GPU_SPAWN_PLACEHOLDER(j): the target of this rank's j-th Commands::Spawn before the ids are assigned."""
    return (np.uint64(SPAWN_PLACEHOLDER_GENERATION) << np.uint64(32)) | np.asarray(j, dtype=np.uint64)


class ApplyStats(C.Structure):
    _fields_ = [
        ("commands", C.c_uint64), ("effective", C.c_uint64), ("invalid", C.c_uint64), ("moved_rows", C.c_uint64),
        ("moved_bytes", C.c_uint64), ("resolve_rounds", C.c_uint32), ("used_sequential_fallback", C.c_uint32),
        ("pipeline", C.c_uint32), ("fast_fallbacks", C.c_uint32),
    ]


class SystemDesc(C.Structure):
    _fields_ = [
        ("name", C.c_char_p), ("query", C.c_void_p), ("launch", C.c_void_p), ("host_user", C.c_void_p),
        ("params", C.c_void_p), ("params_size", C.c_uint32), ("flags", C.c_uint32),
        ("cmd_slots_per_row", C.c_uint32), ("cmd_payload_stride", C.c_uint32),
        ("writes_message", C.c_uint64), ("reads_message", C.c_uint64), ("writer_capacity", C.c_uint64),
        ("n_extra_read_components", C.c_uint32), ("n_extra_write_components", C.c_uint32),
        ("n_read_resources", C.c_uint32), ("n_write_resources", C.c_uint32),
        ("extra_read_components", C.c_uint64 * 8), ("extra_write_components", C.c_uint64 * 8),
        ("read_resources", C.c_uint64 * 8), ("write_resources", C.c_uint64 * 8),
        ("n_more_writes", C.c_uint32), ("n_more_reads", C.c_uint32),
        ("more_writes", C.c_uint64 * 3), ("more_writer_capacity", C.c_uint64 * 3), ("more_reads", C.c_uint64 * 3),
    ]


COND_CUSTOM, COND_TERMINAL, COND_RESOURCE, COND_MESSAGES, COND_CONSTANT, COND_WORLD_RESOURCE = 0, 1, 2, 3, 4, 5
COND_FIELD_VISITED, COND_FIELD_COUNT, COND_FIELD_INDEX, COND_FIELD_I64 = 0, 1, 2, 5
CMP_EQ, CMP_NE, CMP_LT, CMP_LE, CMP_GT, CMP_GE = 0, 1, 2, 3, 4, 5


class ConditionDesc(C.Structure):
    _fields_ = [
        ("name", C.c_char_p), ("kind", C.c_uint32), ("cmp", C.c_uint32), ("rhs", C.c_uint64),
        ("field", C.c_uint32), ("offset", C.c_uint32), ("query", C.c_void_p), ("launch", C.c_void_p),
        ("params", C.c_void_p), ("params_size", C.c_uint32), ("_pad", C.c_uint32), ("reads_message", C.c_uint64),
        ("resource", C.c_uint64),
    ]


class HeliosError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"status {status}: {message}")
        self.status = status


class Backend:
    """One loaded shared library + the symbol prefix of its entry points."""

    def __init__(self, path: str, prefix: str):
        if not os.path.exists(path):
            raise FileNotFoundError(
                f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                f"(there is no CPU fallback for the GPU path)")
        self.lib = C.CDLL(path)
        self.prefix = prefix
        self.path = path
        self.is_gpu = prefix == "helios_gpu_"

    def fn(self, name: str, restype=C.c_int32):
        f = getattr(self.lib, self.prefix + name)
        f.restype = restype
        return f


_GPU = None
_ORACLE = None


def load_gpu() -> Backend:
    global _GPU
    if _GPU is None:
        _GPU = Backend(os.path.join(_HERE, "libhelios_gpu.so"), "helios_gpu_")
    return _GPU


def load_oracle() -> Backend:
    """Test infrastructure only (tests/, smoke(), bench.py cpu_baseline)."""
    global _ORACLE
    if _ORACLE is None:
        _ORACLE = Backend(os.path.join(REPO_ROOT, "oracle", "liboracle.so"), "helios_oracle_")
    return _ORACLE


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class World:
    """helios::ecs::World-shaped handle (reference world.hpp) over one backend."""

    def __init__(self, backend: Backend, device: int = 0, entity_capacity: int = 0, flags: int = 0):
        self.b = backend
        self.h = C.c_void_p()
        cfg = WorldConfig(device, entity_capacity, 0, flags)
        st = backend.fn("world_create")(C.byref(cfg), C.byref(self.h))
        if st != OK:
            raise HeliosError(st, "world_create failed (is a CUDA device visible? there is no CPU fallback)")
        self.slots: dict[str, int] = {}
        self.sparse_names: list[str] = []
        self.sizes: dict[str, int] = {}
        self.hash_to_name: dict[int, str] = {}
        self._keep = []

    # -- plumbing ---------------------------------------------------------------------------------------
    def _check(self, st: int):
        if st != OK:
            msg = self.b.fn("last_error", C.c_char_p)(self.h)
            raise HeliosError(st, msg.decode() if msg else "")

    def close(self):
        if self.h:
            self.b.fn("world_destroy")(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream: int):
        self._check(self.b.fn("world_set_stream")(self.h, C.c_void_p(cuda_stream)))
        self.stream_ptr = cuda_stream

    def require_stream(self, torch, device):
        """Collectives issued through torch run on torch's CURRENT stream; the library's kernels run on the world's.  Code that
        orders one against the other by stream order (the 4-byte all-reduce barriers of the peer-memory exchanges) is only
        correct when the two are the same stream: bind with set_stream(torch.cuda.current_stream(device).cuda_stream)."""
        cur = torch.cuda.current_stream(device).cuda_stream
        if getattr(self, "stream_ptr", None) != cur:
            raise HeliosError(ERR_STATE, "the world's stream is not torch's current stream: call world.set_stream("
                                         "torch.cuda.current_stream(device).cuda_stream) inside the same torch.cuda.stream(...) block")

    def synchronize(self):
        if self.b.is_gpu:
            self._check(self.b.fn("world_synchronize")(self.h))

    def launch_count(self) -> int:
        return int(self.b.fn("world_launch_count", C.c_uint64)(self.h)) if self.b.is_gpu else 0

    # -- components / entities ------------------------------------------------------------------------------
    def register_component(self, name: str, size: int, align: int = 4, default: bytes | None = None) -> int:
        buf = C.create_string_buffer(default, size) if default is not None else None
        slot = self.b.fn("register_component")(self.h, C.c_uint64(fnv1a(name)), C.c_uint32(size), C.c_uint32(align), buf)
        if slot < 0:
            self._check(slot)
        self.slots[name] = slot
        self.sizes[name] = size
        self.hash_to_name[fnv1a(name)] = name
        return slot

    def register_sparse_component(self, name: str, size: int, align: int = 4) -> int:
        """A sparse-set stored component (component/sparse_storage.hpp); size 0 = tag."""
        slot = self.b.fn("register_sparse_component")(self.h, C.c_uint64(fnv1a(name)), C.c_uint32(size), C.c_uint32(align))
        if slot < 0:
            self._check(slot)
        self.slots[name] = slot
        self.sizes[name] = size
        self.hash_to_name[fnv1a(name)] = name
        self.sparse_names.append(name)
        return slot

    def snapshot_sparse(self, name: str) -> dict:
        """Dense order of one sparse storage: reverse map (entity index per dense slot) + raw value bytes."""
        n = C.c_uint64()
        self._check(self.b.fn("snapshot_sparse")(self.h, C.c_uint64(fnv1a(name)), C.byref(n), None, None, C.c_uint64(0)))
        idx = np.empty(n.value, dtype=np.uint32)
        data = np.empty(n.value * self.sizes[name], dtype=np.uint8)
        self._check(self.b.fn("snapshot_sparse")(self.h, C.c_uint64(fnv1a(name)), C.byref(n), _ptr(idx) if idx.size else None,
                                                 _ptr(data) if data.size else None, C.c_uint64(n.value)))
        return {"indices": idx, "data": data}

    def mask(self, names: Iterable[str]) -> int:
        m = 0
        for n in names:
            m |= 1 << self.slots[n]
        return m

    def create_entities(self, n: int) -> np.ndarray:
        out = np.empty(n, dtype=np.uint64)
        self._check(self.b.fn("create_entities")(self.h, C.c_uint32(n), _ptr(out)))
        return out

    def reserve_entities(self, n: int) -> np.ndarray:
        out = np.empty(n, dtype=np.uint64)
        self._check(self.b.fn("reserve_entities")(self.h, C.c_uint32(n), _ptr(out)))
        return out

    def spawn_batch(self, names: Sequence[str], n: int, columns: Sequence[np.ndarray], want_entities: bool = True):
        hashes = (C.c_uint64 * max(1, len(names)))(*[fnv1a(x) for x in names])
        cols = [np.ascontiguousarray(c) for c in columns]
        for nm, c in zip(names, cols):
            assert c.nbytes == n * self.sizes[nm], (nm, c.nbytes, n, self.sizes[nm])
        ptrs = (C.c_void_p * max(1, len(names)))(*[c.ctypes.data for c in cols])
        out = np.empty(n, dtype=np.uint64) if want_entities else None
        self._check(self.b.fn("spawn_batch")(self.h, hashes, C.c_uint32(len(names)), C.c_uint64(n), ptrs,
                                             _ptr(out) if out is not None else None))
        return out

    def clear_entities(self):
        """World::ClearEntities (world.hpp:974-977)."""
        self._check(self.b.fn("world_clear_entities")(self.h))

    def extract_from(self, query: "Query") -> int:
        """SubApp::Extract with the usual body: one entity here per row `query` (of another world) visits, carrying copies
        of its data components, in visit order.  Returns the number of entities created."""
        for name in query.component_names:  # keep this wrapper's name tables in step (the library registers on demand too)
            if name not in self.slots:
                self.register_component(name, query.w.sizes[name])
        n = C.c_uint64()
        if self.b.is_gpu:
            self._check(self.b.fn("extract_query")(query.h, self.h, C.byref(n)))
        else:
            self._check(self.b.fn("extract_query_w")(query.w.h, query.h, self.h, C.byref(n)))
        return n.value

    def entity_count(self) -> int:
        v = C.c_uint64()
        self._check(self.b.fn("entity_count")(self.h, C.byref(v)))
        return v.value

    def entity_location(self, e: int):
        a, r = C.c_uint32(), C.c_uint32()
        st = self.b.fn("entity_location")(self.h, C.c_uint64(int(e)), C.byref(a), C.byref(r))
        if st == ERR_NOT_FOUND:
            return None
        self._check(st)
        return a.value, r.value

    # -- snapshot ---------------------------------------------------------------------------------------------
    def snapshot(self, columns: bool = True) -> dict:
        """Canonical dump: archetypes in creation order (sorted type hashes, entities in row order, raw column bytes)
        plus the entity tables."""
        n = C.c_uint32()
        self._check(self.b.fn("snapshot_archetype_count")(self.h, C.byref(n)))
        archs = []
        for a in range(n.value):
            hashes = (C.c_uint64 * MAX_COMPONENTS)()
            nt, rows = C.c_uint32(), C.c_uint64()
            self._check(self.b.fn("snapshot_archetype_info")(self.h, C.c_uint32(a), hashes, C.byref(nt), C.byref(rows)))
            types = [int(hashes[i]) for i in range(nt.value)]
            ents = np.empty(rows.value, dtype=np.uint64)
            self._check(self.b.fn("snapshot_archetype_entities")(self.h, C.c_uint32(a), _ptr(ents), C.c_uint64(rows.value)))
            cols = {}
            if columns:
                for t in types:
                    size = self.sizes[self.hash_to_name[t]]
                    buf = np.empty(rows.value * size, dtype=np.uint8)
                    self._check(self.b.fn("snapshot_archetype_column")(self.h, C.c_uint32(a), C.c_uint64(t), _ptr(buf), C.c_uint64(buf.nbytes)))
                    cols[t] = buf
            archs.append({"types": types, "entities": ents, "columns": cols})
        nxt, fc, cur = C.c_uint32(), C.c_uint32(), C.c_int64()
        self._check(self.b.fn("snapshot_entity_table")(self.h, C.byref(nxt), None, C.c_uint32(0), None, C.c_uint32(0), C.byref(fc), C.byref(cur)))
        gens = np.full(nxt.value, 0xFFFFFFFF, dtype=np.uint32)
        free = np.empty(fc.value, dtype=np.uint32)
        self._check(self.b.fn("snapshot_entity_table")(self.h, C.byref(nxt), _ptr(gens), C.c_uint32(gens.size), _ptr(free), C.c_uint32(free.size), C.byref(fc), C.byref(cur)))
        return {"archetypes": archs, "next_index": nxt.value, "generations": gens, "free_indices": free,
                "free_cursor": cur.value, "entity_count": self.entity_count(), "messages": self.snapshot_messages()}

    # -- message queues as part of the world state (builtin structural messages included) ----------------------------
    def message_types(self) -> list[tuple[int, int]]:
        """(type hash, element size) of every registered non-async message type, sorted by hash."""
        n = C.c_uint32()
        self._check(self.b.fn("message_types")(self.h, None, None, C.c_uint32(0), C.byref(n)))
        hashes = (C.c_uint64 * max(1, n.value))()
        sizes = (C.c_uint32 * max(1, n.value))()
        self._check(self.b.fn("message_types")(self.h, hashes, sizes, C.c_uint32(n.value), C.byref(n)))
        return sorted((int(hashes[i]), int(sizes[i])) for i in range(n.value))

    def message_read_hash(self, type_hash: int, size: int, which: int) -> np.ndarray:
        p, c = C.c_uint64(), C.c_uint64()
        self._check(self.b.fn("message_counts")(self.h, C.c_uint64(type_hash), C.byref(p), C.byref(c)))
        cnt = c.value if which else p.value
        buf = np.empty(cnt * size, dtype=np.uint8)
        self._check(self.b.fn("message_read")(self.h, C.c_uint64(type_hash), C.c_uint32(which), _ptr(buf) if buf.size else None,
                                              C.c_uint64(buf.nbytes)))
        return buf

    def snapshot_messages(self) -> dict:
        """{type hash: (previous bytes, current bytes)} over every registered message type."""
        return {h: (self.message_read_hash(h, sz, 0), self.message_read_hash(h, sz, 1)) for h, sz in self.message_types()}

    def builtin_message_type(self, kind: int, component: str | None = None) -> int:
        out = C.c_uint64()
        self._check(self.b.fn("builtin_message_type")(self.h, C.c_uint32(kind), C.c_uint64(fnv1a(component) if component else 0),
                                                      C.byref(out)))
        return out.value

    def component_message_types(self, component: str, added_hash: int, removed_hash: int):
        self._check(self.b.fn("component_message_types")(self.h, C.c_uint64(fnv1a(component)), C.c_uint64(added_hash),
                                                         C.c_uint64(removed_hash)))

    def builtin_messages(self, kind: int, component: str | None = None, which: int = 1) -> np.ndarray:
        """Entities carried by the builtin messages of one kind (previous: which = 0, current: which = 1); empty if the
        type was never registered."""
        h = self.builtin_message_type(kind, component)
        if h not in dict(self.message_types()):
            return np.empty(0, dtype=np.uint64)
        return self.message_read_hash(h, 8, which).view(np.uint64)

    def column_write(self, archetype: int, name: str, row0: int, rows: int, src_ptr: int):
        self._check(self.b.fn("archetype_column_write")(self.h, C.c_uint32(archetype), C.c_uint64(fnv1a(name)),
                                                        C.c_uint64(row0), C.c_uint64(rows), C.c_void_p(src_ptr)))

    def column_write_on(self, stream_ptr: int, archetype: int, name: str, row0: int, rows: int, src_ptr: int):
        self._check(self.b.fn("archetype_column_write_on")(self.h, C.c_uint32(archetype), C.c_uint64(fnv1a(name)),
                                                           C.c_uint64(row0), C.c_uint64(rows), C.c_void_p(src_ptr),
                                                           C.c_void_p(stream_ptr)))

    def column_read(self, archetype: int, name: str, row0: int, rows: int, dst_ptr: int):
        self._check(self.b.fn("archetype_column_read")(self.h, C.c_uint32(archetype), C.c_uint64(fnv1a(name)),
                                                       C.c_uint64(row0), C.c_uint64(rows), C.c_void_p(dst_ptr)))

    # -- queries ----------------------------------------------------------------------------------------------
    def query(self, components: Sequence[tuple[str, bool]], with_: Sequence[str] = (), without: Sequence[str] = (),
              sparse_data: Sequence[str] = (), sparse_rule: int = SPARSE_STRICT,
              optional: Sequence[tuple[str, bool]] = ()) -> "Query":
        """optional: (component, write) of `T*` / `const T*` arguments over archetype-stored components."""
        return Query(self, components, with_, without, sparse_data, sparse_rule, optional)

    # -- commands ---------------------------------------------------------------------------------------------
    def enqueue(self, cmds: np.ndarray, payload: bytes | np.ndarray = b""):
        cmds = np.ascontiguousarray(cmds, dtype=COMMAND_DTYPE)
        pay = np.frombuffer(payload, dtype=np.uint8) if isinstance(payload, (bytes, bytearray)) else np.ascontiguousarray(payload).view(np.uint8)
        self._check(self.b.fn("commands_enqueue")(self.h, _ptr(cmds), C.c_uint64(cmds.size), _ptr(pay) if pay.size else None, C.c_uint64(pay.size)))

    def stage_commands(self):
        if self.b.is_gpu:
            self._check(self.b.fn("commands_stage")(self.h))

    def flush(self):
        self._check(self.b.fn("world_flush")(self.h))

    def next_entity_index(self) -> int:
        """EntityManager::next_index_ (entity/manager.hpp:294-305): the index the next fresh reservation takes."""
        v = C.c_uint32()
        self._check(self.b.fn("snapshot_entity_table")(self.h, C.byref(v), None, C.c_uint32(0), None, C.c_uint32(0), None, None))
        return v.value

    def commands_route_pack(self, cmds: np.ndarray, payload: bytes | np.ndarray, spawn_first_index: int) -> np.ndarray:
        """Queue records + payload -> 64-byte routed records (spawn placeholders become global ids)."""
        cmds = np.ascontiguousarray(cmds, dtype=COMMAND_DTYPE)
        pay = np.frombuffer(payload, dtype=np.uint8) if isinstance(payload, (bytes, bytearray)) else np.ascontiguousarray(payload).view(np.uint8)
        out = np.zeros(cmds.size, dtype=ROUTED_COMMAND_DTYPE)
        self._check(self.b.fn("commands_route_pack")(self.h, _ptr(cmds) if cmds.size else None, C.c_uint64(cmds.size),
                                                     _ptr(pay) if pay.size else None, C.c_uint64(pay.size),
                                                     C.c_uint32(spawn_first_index), _ptr(out) if cmds.size else None))
        return out

    def commands_apply_routed(self, name: str, which: int = 1):
        self._check(self.b.fn("commands_apply_routed")(self.h, C.c_uint64(fnv1a(name)), C.c_uint32(which)))

    def update(self):
        self._check(self.b.fn("world_update")(self.h))

    def apply_stats(self) -> ApplyStats:
        s = ApplyStats()
        self._check(self.b.fn("last_apply_stats")(self.h, C.byref(s)))
        return s

    # -- messages ---------------------------------------------------------------------------------------------
    def message_register(self, name: str, size: int, policy: int = MSG_AUTOMATIC):
        self.sizes["msg:" + name] = size
        self._check(self.b.fn("message_register")(self.h, C.c_uint64(fnv1a(name)), C.c_uint32(size), C.c_uint32(policy)))

    def message_write(self, name: str, data: np.ndarray):
        data = np.ascontiguousarray(data)
        n = data.nbytes // self.sizes["msg:" + name]
        self._check(self.b.fn("message_write")(self.h, C.c_uint64(fnv1a(name)), _ptr(data), C.c_uint64(n)))

    def message_counts(self, name: str):
        p, c = C.c_uint64(), C.c_uint64()
        self._check(self.b.fn("message_counts")(self.h, C.c_uint64(fnv1a(name)), C.byref(p), C.byref(c)))
        return p.value, c.value

    def message_read(self, name: str, which: int) -> np.ndarray:
        cnt = self.message_counts(name)[which]
        buf = np.empty(cnt * self.sizes["msg:" + name], dtype=np.uint8)
        self._check(self.b.fn("message_read")(self.h, C.c_uint64(fnv1a(name)), C.c_uint32(which), _ptr(buf), C.c_uint64(buf.nbytes)))
        return buf

    # ---- world resources (reference resource/manager.hpp; world.hpp:491-620) ----
    def resource_insert(self, name: str, value: bytes):
        buf = C.create_string_buffer(value, len(value))
        self._check(self.b.fn("resource_insert")(self.h, C.c_uint64(fnv1a(name)), buf, C.c_uint32(len(value))))
        self.sizes["res:" + name] = len(value)

    def resource_try_insert(self, name: str, value: bytes) -> bool:
        buf = C.create_string_buffer(value, len(value))
        ok = C.c_uint32()
        self._check(self.b.fn("resource_try_insert")(self.h, C.c_uint64(fnv1a(name)), buf, C.c_uint32(len(value)), C.byref(ok)))
        if ok.value:
            self.sizes["res:" + name] = len(value)
        return bool(ok.value)

    def resource_remove(self, name: str):
        self._check(self.b.fn("resource_remove")(self.h, C.c_uint64(fnv1a(name))))

    def resource_try_remove(self, name: str) -> bool:
        ok = C.c_uint32()
        self._check(self.b.fn("resource_try_remove")(self.h, C.c_uint64(fnv1a(name)), C.byref(ok)))
        return bool(ok.value)

    def resource_has(self, name: str) -> bool:
        v = C.c_uint32()
        self._check(self.b.fn("resource_has")(self.h, C.c_uint64(fnv1a(name)), C.byref(v)))
        return bool(v.value)

    def resource_count(self) -> int:
        v = C.c_uint32()
        self._check(self.b.fn("resource_count")(self.h, C.byref(v)))
        return v.value

    def resource_clear(self):
        self._check(self.b.fn("resource_clear")(self.h))

    def resource_read(self, name: str) -> bytes:
        n = self.sizes["res:" + name]
        buf = C.create_string_buffer(n)
        self._check(self.b.fn("resource_read")(self.h, C.c_uint64(fnv1a(name)), buf, C.c_uint32(n)))
        return buf.raw

    def resource_write(self, name: str, value: bytes):
        buf = C.create_string_buffer(value, len(value))
        self._check(self.b.fn("resource_write")(self.h, C.c_uint64(fnv1a(name)), buf, C.c_uint32(len(value))))

    # ---- async messages (reference message/async_queue.hpp; MessageManager::WriteAsync / ManualAsyncClear) ----
    def async_message_register(self, name: str, size: int, capacity: int = 1024):
        self._check(self.b.fn("async_message_register")(self.h, C.c_uint64(fnv1a(name)), C.c_uint32(size), C.c_uint64(capacity)))
        self.sizes["async:" + name] = size

    def async_message_write(self, name: str, data: np.ndarray):
        size = self.sizes["async:" + name]
        buf = np.ascontiguousarray(data).view(np.uint8).reshape(-1)
        self._check(self.b.fn("async_message_write")(self.h, C.c_uint64(fnv1a(name)), _ptr(buf) if buf.size else None,
                                                      C.c_uint64(buf.size // size)))

    def async_message_dequeue(self, name: str, max_count: int = 1 << 62) -> np.ndarray:
        """AsyncMessageReader<T>::Into(out, max_count): the dequeued messages as raw bytes, queue order."""
        size = self.sizes["async:" + name]
        want = min(max_count, self.async_message_size(name))
        buf = np.zeros(max(1, want) * size, dtype=np.uint8)
        n = C.c_uint64()
        self._check(self.b.fn("async_message_dequeue")(self.h, C.c_uint64(fnv1a(name)), _ptr(buf), C.c_uint64(want), C.byref(n)))
        return buf[: n.value * size]

    def async_message_size(self, name: str) -> int:
        n = C.c_uint64()
        self._check(self.b.fn("async_message_size")(self.h, C.c_uint64(fnv1a(name)), C.byref(n)))
        return n.value

    def async_message_clear(self, name: str | None = None):
        self._check(self.b.fn("async_message_clear")(self.h, C.c_uint64(fnv1a(name) if name else 0)))

    def schedule(self, name: str = "Update") -> "Schedule":
        return Schedule(self, name)


class Query:
    """Query<Cs..., With<...>, Without<...>> (reference query/query.hpp:658-1248)."""

    def __init__(self, world: World, components, with_, without, sparse_data=(), sparse_rule=SPARSE_STRICT, optional=()):
        self.w = world
        self.component_names = [name for name, _ in components]
        d = QueryDesc()
        d.n_optional = len(optional)
        for i, (name, write) in enumerate(optional):
            d.optional[i] = fnv1a(name)
            d.optional_writes[i] = 1 if write else 0
        d.n_sparse_data = len(sparse_data)
        for i, name in enumerate(sparse_data):
            d.sparse_data[i] = fnv1a(name)
        d.sparse_rule = sparse_rule
        d.n_components = len(components)
        for i, (name, write) in enumerate(components):
            d.components[i] = fnv1a(name)
            d.writes[i] = 1 if write else 0
        d.n_with = len(with_)
        for i, name in enumerate(with_):
            d.with_[i] = fnv1a(name)
        d.n_without = len(without)
        for i, name in enumerate(without):
            d.without[i] = fnv1a(name)
        self.h = C.c_void_p()
        world._check(world.b.fn("query_create")(world.h, C.byref(d), C.byref(self.h)))

    def match(self) -> list[int]:
        out = (C.c_uint32 * 4096)()
        n = C.c_uint32()
        if self.w.b.is_gpu:
            self.w._check(self.w.b.fn("query_match")(self.h, out, C.c_uint32(4096), C.byref(n)))
        else:
            self.w._check(self.w.b.fn("query_match_w")(self.w.h, self.h, out, C.c_uint32(4096), C.byref(n)))
        return [out[i] for i in range(n.value)]

    def count(self) -> int:
        v = C.c_uint64()
        if self.w.b.is_gpu:
            self.w._check(self.w.b.fn("query_count")(self.h, C.byref(v)))
        else:
            self.w._check(self.w.b.fn("query_count_w")(self.w.h, self.h, C.byref(v)))
        return v.value


    # -- terminals (query.hpp:180,822,1040-1180) -----------------------------------------------------------------------------
    def terminal(self, name: str, params: bytes = b"") -> "TerminalResult":
        """Runs a built-in terminal (helios_gpu_builtin_terminal) over this query and returns its result."""
        b = self.w.b
        fn, flags = C.c_void_p(), C.c_uint32()
        self.w._check(b.fn("builtin_terminal")(name.encode(), C.byref(fn), C.byref(flags)))
        out = TerminalResult()
        buf = C.create_string_buffer(params, len(params)) if params else None
        if b.is_gpu:
            self.w._check(b.fn("query_terminal")(self.h, fn, buf, C.c_uint32(len(params)), flags, C.byref(out)))
        else:
            self.w._check(b.fn("query_terminal_w")(self.w.h, self.h, fn, buf, C.c_uint32(len(params)), flags, C.byref(out)))
        return out

    def terminal_entities(self, group: int) -> np.ndarray:
        """Entities of the last partition-style terminal: group 0 = predicate true, 1 = the rest (iteration order)."""
        n = C.c_uint64()
        rc = self.w.b.fn("query_terminal_entities")(self.h, C.c_uint32(group), None, C.c_uint64(0), C.byref(n))
        if rc not in (0, ERR_CAPACITY):
            self.w._check(rc)
        out = np.empty(n.value, dtype=np.uint64)
        if n.value:
            self.w._check(self.w.b.fn("query_terminal_entities")(self.h, C.c_uint32(group), out.ctypes.data_as(C.c_void_p),
                                                                 C.c_uint64(n.value), C.byref(n)))
        return out

    # reference spellings on top of the built-in predicate `Position.x > t`
    def count_if_pos_x_gt(self, t: float) -> int:
        return self.terminal("count_if_pos_x_gt", struct.pack("<f", t)).count

    def any_pos_x_gt(self, t: float) -> bool:  # Any = Find(...).has_value()  (query.hpp:1142-1144)
        return self.terminal("find_pos_x_gt", struct.pack("<f", t)).index != NO_ROW

    def none_pos_x_gt(self, t: float) -> bool:  # None = !Any  (query.hpp:1166-1168)
        return not self.any_pos_x_gt(t)

    def all_pos_x_gt(self, t: float) -> bool:  # All: no row fails; vacuously true on an empty match set
        r = self.terminal("count_if_pos_x_gt", struct.pack("<f", t))
        return r.count == r.visited


NO_ROW = 0xFFFFFFFFFFFFFFFF


class TerminalResult(C.Structure):
    _fields_ = [("visited", C.c_uint64), ("count", C.c_uint64), ("index", C.c_uint64), ("entity", C.c_uint64),
                ("f64", C.c_double), ("i64", C.c_int64)]


class Schedule:
    """helios::ecs::Schedule (reference schedule/schedule.hpp, src/ecs/src/schedule/schedule.cpp)."""

    def __init__(self, world: World, name: str):
        self.w = world
        self.h = C.c_void_p()
        world._check(world.b.fn("schedule_create")(world.h, name.encode(), C.byref(self.h)))
        self._keep = []
        self.names: list[str] = []

    def add_system(self, name: str, builtin: str, query: Query | None = None, params: bytes = b"", flags: int = 0,
                   writes_message: str | None = None, reads_message: str | None = None, writer_capacity: int = 0,
                   cmd_slots_per_row: int = 1, cmd_payload_stride: int = 0,
                   extra_reads: Sequence[str] = (), extra_writes: Sequence[str] = (),
                   read_resources: Sequence[str] = (), write_resources: Sequence[str] = (),
                   more_writes: Sequence[tuple[str, int]] = (), more_reads: Sequence[str] = ()) -> int:
        """more_writes: (message type, writer capacity) of the 2nd.. MessageWriter parameter; more_reads: 2nd.. reader."""
        fnp = C.c_void_p()
        st = self.w.b.fn("builtin_system")(builtin.encode(), C.byref(fnp))
        if st != OK:
            raise HeliosError(st, f"no builtin system '{builtin}'")
        d = SystemDesc()
        nm = name.encode()
        pb = C.create_string_buffer(params, len(params)) if params else None
        self._keep += [nm, pb]
        d.name = nm
        d.query = query.h if query is not None else None
        d.launch = fnp
        d.params = C.cast(pb, C.c_void_p) if pb is not None else None
        d.params_size = len(params)
        d.flags = flags
        d.cmd_slots_per_row = cmd_slots_per_row
        d.cmd_payload_stride = cmd_payload_stride
        d.writes_message = fnv1a(writes_message) if writes_message else 0
        d.reads_message = fnv1a(reads_message) if reads_message else 0
        d.writer_capacity = writer_capacity
        d.n_more_writes = len(more_writes)
        for i, (mname, mcap) in enumerate(more_writes):
            d.more_writes[i] = fnv1a(mname)
            d.more_writer_capacity[i] = mcap
        d.n_more_reads = len(more_reads)
        for i, mname in enumerate(more_reads):
            d.more_reads[i] = fnv1a(mname)
        for attr, cnt, names in (("extra_read_components", "n_extra_read_components", extra_reads),
                                 ("extra_write_components", "n_extra_write_components", extra_writes),
                                 ("read_resources", "n_read_resources", read_resources),
                                 ("write_resources", "n_write_resources", write_resources)):
            setattr(d, cnt, len(names))
            arr = getattr(d, attr)
            for i, x in enumerate(names):
                arr[i] = fnv1a(x)
        idx = C.c_uint32()
        self.w._check(self.w.b.fn("schedule_add_system")(self.h, C.byref(d), C.byref(idx)))
        self.names.append(name)
        return idx.value

    def _condition(self, name: str, kind: int, cmp: int = CMP_NE, rhs: int = 0, field: int = 0, offset: int = 0,
                   query: "Query | None" = None, builtin: str | None = None, params: bytes = b"",
                   reads_message: str | None = None, resource: str | None = None) -> ConditionDesc:
        d = ConditionDesc()
        nm = name.encode()
        pb = C.create_string_buffer(params, len(params)) if params else None
        self._keep += [nm, pb]
        d.name, d.kind, d.cmp, d.rhs, d.field, d.offset = nm, kind, cmp, rhs & 0xFFFFFFFFFFFFFFFF, field, offset
        d.query = query.h if query is not None else None
        if builtin is not None:
            fnp = C.c_void_p()
            st = self.w.b.fn("builtin_condition")(builtin.encode(), C.byref(fnp))
            if st != OK:
                raise HeliosError(st, f"no builtin condition '{builtin}'")
            d.launch = fnp
        d.params = C.cast(pb, C.c_void_p) if pb is not None else None
        d.params_size = len(params)
        d.reads_message = fnv1a(reads_message) if reads_message else 0
        d.resource = fnv1a(resource) if resource else 0
        return d

    def run_if(self, system: int, name: str, kind: int, **kw):
        """SystemHandle::RunIf (reference schedule/system_handle.hpp; run_condition.hpp:23-58)."""
        d = self._condition(name, kind, **kw)
        self.w._check(self.w.b.fn("schedule_add_run_condition")(self.h, C.c_uint32(system), C.byref(d)))

    def set_run_if(self, set_id: int, name: str, kind: int, **kw):
        """SystemSet::RunIf (reference schedule/system_set.hpp)."""
        d = self._condition(name, kind, **kw)
        self.w._check(self.w.b.fn("schedule_add_set_condition")(self.h, C.c_uint64(set_id), C.byref(d)))

    def in_set(self, system: int, set_id: int):
        self.w._check(self.w.b.fn("schedule_system_in_set")(self.h, C.c_uint32(system), C.c_uint64(set_id)))

    def last_verdicts(self) -> list[int]:
        out = (C.c_uint32 * 64)()
        self.w._check(self.w.b.fn("schedule_last_verdicts")(self.h, out, C.c_uint32(64)))
        return [int(out[i]) for i in range(len(self.names))]

    # -- system-set ordering (system_set_handle.hpp:42-230, system_handle.hpp:105-180; schedule.cpp:209-339) --
    def ensure_set(self, set_id: int):
        """Schedule::Set(id)."""
        self.w._check(self.w.b.fn("schedule_ensure_set")(self.h, C.c_uint64(set_id)))

    def order_set_system(self, set_id: int, system: int, after: bool = False):
        """Set(set_id).Before(system) / .After(system)."""
        self.w._check(self.w.b.fn("schedule_order_set_system")(self.h, C.c_uint64(set_id), C.c_uint32(system), C.c_uint32(int(after))))

    def order_sets(self, set_id: int, target_set: int, after: bool = False):
        """Set(set_id).Before(SystemSetId) / .After(SystemSetId) — BeforeSet<T> / AfterSet<T>."""
        self.w._check(self.w.b.fn("schedule_order_sets")(self.h, C.c_uint64(set_id), C.c_uint64(target_set), C.c_uint32(int(after))))

    def order_system_set(self, system: int, set_id: int, after: bool = False):
        """SystemHandle::Before(SystemSetId) / After(SystemSetId)."""
        self.w._check(self.w.b.fn("schedule_order_system_set")(self.h, C.c_uint32(system), C.c_uint64(set_id), C.c_uint32(int(after))))

    def set_sequence(self, set_id: int):
        """Set(set_id).Sequence()."""
        self.w._check(self.w.b.fn("schedule_set_sequence")(self.h, C.c_uint64(set_id)))

    def add_order(self, first: int, second: int):
        self.w._check(self.w.b.fn("schedule_add_order")(self.h, C.c_uint32(first), C.c_uint32(second)))

    def build(self):
        self.w._check(self.w.b.fn("schedule_build")(self.h))

    def execution_order(self) -> list[int]:
        out = (C.c_uint32 * 64)()
        n = C.c_uint32()
        self.w._check(self.w.b.fn("schedule_execution_order")(self.h, out, C.c_uint32(64), C.byref(n)))
        return [out[i] for i in range(n.value)]

    def conflicts(self) -> list[int]:
        n = len(self.names)
        out = (C.c_uint64 * 64)()
        self.w._check(self.w.b.fn("schedule_conflicts")(self.h, out, C.c_uint32(64)))
        return [int(out[i]) for i in range(n)]

    def stream_count(self) -> int:
        v = C.c_uint32()
        self.w._check(self.w.b.fn("schedule_stream_count")(self.h, C.byref(v)))
        return v.value

    def run(self, mode: int = EXEC_SERIAL):
        self.w._check(self.w.b.fn("schedule_run")(self.h, C.c_uint32(mode)))

    def apply_deferred(self):
        self.w._check(self.w.b.fn("schedule_apply_deferred")(self.h))

    def run_and_wait(self, mode: int = EXEC_SERIAL):
        self.w._check(self.w.b.fn("schedule_run_and_wait")(self.h, C.c_uint32(mode)))

    def read_resources(self, nbytes: int = 64) -> np.ndarray:
        buf = np.zeros(nbytes, dtype=np.uint8)
        self.w._check(self.w.b.fn("schedule_read_resources")(self.h, _ptr(buf), C.c_uint32(nbytes)))
        return buf

    def close(self):
        if self.h:
            self.w.b.fn("schedule_destroy")(self.h)
            self.h = C.c_void_p()
