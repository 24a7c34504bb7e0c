"""GSD files (General Simulation Data, file format 1.0) with the HOOMD schema 1.2 — what `hoomd.dump.gsd`,
`hoomd.init.read_gsd` and `hoomd.data.gsd_snapshot` read and write (snapshot / restart I/O on the
device-resident state).

File layer = an independent implementation of the container the reference vendors as hoomd/extern/gsd.c (layout:
hoomd/extern/gsd.h:44-86): a 256-byte header, an index block of 32-byte entries {frame, N, location, M, id, type, flags},
a namelist of 64-byte names, data chunks appended at the end of the file, the index doubled and moved to the end when it
is full.  The write path follows gsd.c's APPEND-mode decisions (hoomd/extern/gsd.c:67-147, 693-808), so a file written here
is byte-identical to one gsd.c writes for the same sequence of chunks; tests/test_gsd.py checks exactly that against the
reference's own gsd.c (compiled into oracle/_ref) and reads each side's files with the other.

Schema layer = GSDDumpWriter::analyze / GSDReader (hoomd/GSDDumpWriter.cc:160-345, 350-726, 727-846; hoomd/GSDReader.cc:
129-371): which chunks a frame holds, the defaults that are NOT written, frame-0 fallback when reading.

Pure numpy: file I/O is not on the hot path; the particle arrays come from / go to the engine's host-buffer calls."""
import os
import struct

import numpy as np

MAGIC = 0x65DF65DF65DF65DF
HEADER = struct.Struct("<QQQQQII64s64s80s")        # gsd_header, 256 bytes
ENTRY = np.dtype([("frame", "<u8"), ("N", "<u8"), ("location", "<i8"), ("M", "<u4"), ("id", "<u2"), ("type", "u1"),
                  ("flags", "u1")])                   # gsd_index_entry, 32 bytes
NAME_BYTES = 64
INITIAL_INDEX = 128
INITIAL_NAMES = 128

# enum gsd_type (hoomd/extern/gsd.h:20-32)
TYPES = {1: np.uint8, 2: np.uint16, 3: np.uint32, 4: np.uint64, 5: np.int8, 6: np.int16, 7: np.int32, 8: np.int64,
         9: np.float32, 10: np.float64}
TYPE_OF = {np.dtype(v): k for k, v in TYPES.items()}


def make_version(major, minor):
    return (major << 16) | minor


def create(path, application, schema="hoomd", schema_version=(1, 2)):
    """gsd_create: an empty file = header + zeroed index (128 entries) + zeroed namelist (128 names)"""
    index_location = HEADER.size
    namelist_location = index_location + ENTRY.itemsize * INITIAL_INDEX
    hdr = HEADER.pack(MAGIC, index_location, INITIAL_INDEX, namelist_location, INITIAL_NAMES,
                      make_version(*schema_version), make_version(1, 0), application.encode()[:63], schema.encode()[:63],
                      b"")
    with open(path, "wb") as f:
        f.write(hdr)
        f.write(bytes(ENTRY.itemsize * INITIAL_INDEX))
        f.write(bytes(NAME_BYTES * INITIAL_NAMES))


class GSDFile:
    """An open GSD file; mode 'r' (read only) or 'a' (append frames, as GSDDumpWriter opens it)."""

    def __init__(self, path, mode="r"):
        if mode not in ("r", "a"):
            raise ValueError("mode must be 'r' or 'a'")
        self.path, self.mode = path, mode
        self.f = open(path, "rb" if mode == "r" else "r+b")
        self._read_header()

    # ---- __gsd_read_header (hoomd/extern/gsd.c:262-420) ----
    def _read_header(self):
        f = self.f
        f.seek(0)
        raw = f.read(HEADER.size)
        if len(raw) != HEADER.size:
            raise RuntimeError("%s is not a valid GSD file" % self.path)
        (magic, self.index_location, self.index_allocated, self.namelist_location, self.namelist_allocated,
         self.schema_version, self.gsd_version, app, schema, _res) = HEADER.unpack(raw)
        if magic != MAGIC:
            raise RuntimeError("%s is not a valid GSD file" % self.path)
        if (self.gsd_version < make_version(1, 0) and self.gsd_version != make_version(0, 3)) \
                or self.gsd_version >= make_version(2, 0):
            raise RuntimeError("Invalid GSD file version in %s" % self.path)
        self.application = app.split(b"\0")[0].decode()
        self.schema = schema.split(b"\0")[0].decode()
        f.seek(0, os.SEEK_END)
        self.file_size = f.tell()
        if self.index_location + ENTRY.itemsize * self.index_allocated > self.file_size \
                or self.namelist_location + NAME_BYTES * self.namelist_allocated > self.file_size:
            raise RuntimeError("Corrupt GSD file: %s" % self.path)
        f.seek(self.index_location)
        index = np.frombuffer(f.read(ENTRY.itemsize * self.index_allocated), ENTRY)
        used = index["location"] != 0
        self.index_num_entries = int(np.argmin(used)) if not used.all() else len(index)
        if self.index_num_entries and not used[:self.index_num_entries].all():
            raise RuntimeError("Corrupt GSD file: %s" % self.path)
        self.index = index[:self.index_num_entries].copy()
        f.seek(self.namelist_location)
        raw = f.read(NAME_BYTES * self.namelist_allocated)
        self.names = []
        for k in range(self.namelist_allocated):
            name = raw[k * NAME_BYTES:(k + 1) * NAME_BYTES].split(b"\0")[0]
            if not name:
                break
            self.names.append(name.decode())
        self.cur_frame = int(self.index["frame"][-1]) + 1 if self.index_num_entries else 0
        self.index_written_entries = self.index_num_entries
        self._pending = []   # index entries of the frame being written (APPEND mode keeps only these in memory)
        self._lookup = {(int(e["frame"]), int(e["id"])): k for k, e in enumerate(self.index)}

    @property
    def nframes(self):
        return self.cur_frame

    def close(self):
        if self.f:
            self.f.close()
            self.f = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- writing (gsd_write_chunk / gsd_end_frame / __gsd_expand_index / gsd_truncate) ----
    def _write_header(self):
        self.f.seek(0)
        self.f.write(HEADER.pack(MAGIC, self.index_location, self.index_allocated, self.namelist_location,
                                 self.namelist_allocated, self.schema_version, self.gsd_version,
                                 self.application.encode()[:63], self.schema.encode()[:63], b""))

    def _name_id(self, name, append):
        if name in self.names:
            return self.names.index(name)
        if not append or len(self.names) >= self.namelist_allocated:
            return 0xFFFF
        enc = name.encode()[:NAME_BYTES - 1]
        self.f.seek(self.namelist_location + NAME_BYTES * len(self.names))
        self.f.write(enc + bytes(NAME_BYTES - len(enc)))
        self.names.append(enc.decode())
        return len(self.names) - 1

    def _expand_index(self):
        old = self.index_allocated
        self.index_allocated = old * 2
        # the index as it is ON DISK (entries of the frame in flight are not there yet) moves to the end of the file
        self.f.seek(self.index_location)
        disk = self.f.read(ENTRY.itemsize * old)
        self.f.seek(0, os.SEEK_END)
        new_location = self.f.tell()
        self.f.write(disk)
        self.f.write(bytes(ENTRY.itemsize * old))
        self.index_location = new_location
        self.file_size = new_location + ENTRY.itemsize * self.index_allocated
        self._write_header()

    def write_chunk(self, name, data):
        if self.mode != "a":
            raise RuntimeError("file is opened read-only")
        a = np.ascontiguousarray(data)
        if a.ndim == 0:
            a = a.reshape(1, 1)
        elif a.ndim == 1:
            a = a.reshape(-1, 1)
        elif a.ndim != 2:
            raise ValueError("chunks are N x M arrays")
        if a.dtype not in TYPE_OF:
            raise ValueError("no GSD type for dtype %s" % a.dtype)
        if a.shape[1] == 0:
            raise ValueError("M must be positive")
        e = np.zeros(1, ENTRY)
        e["frame"], e["N"], e["M"] = self.cur_frame, a.shape[0], a.shape[1]
        e["id"] = self._name_id(name, True)
        e["type"] = TYPE_OF[a.dtype]
        e["location"] = self.file_size
        self.f.seek(self.file_size)
        self.f.write(a.astype(a.dtype.newbyteorder("<"), copy=False).tobytes())
        self.file_size += a.nbytes
        if self.index_num_entries >= self.index_allocated:
            self._expand_index()
        self._pending.append(e)
        self.index_num_entries += 1

    def end_frame(self):
        if self.mode != "a":
            raise RuntimeError("file is opened read-only")
        self.cur_frame += 1
        if self._pending:
            block = np.concatenate(self._pending)
            self.f.seek(self.index_location + ENTRY.itemsize * self.index_written_entries)
            self.f.write(block.tobytes())
            base = len(self.index)
            self.index = np.concatenate([self.index, block])
            for k, e in enumerate(block):
                self._lookup[(int(e["frame"]), int(e["id"]))] = base + k
            self.index_written_entries += len(block)
            self._pending = []
        self.f.flush()

    def truncate(self):
        """gsd_truncate: back to an empty file with the same application / schema"""
        if self.mode != "a":
            raise RuntimeError("file is opened read-only")
        app, schema, ver = self.application, self.schema, self.schema_version
        self.f.close()
        create(self.path, app, schema, (ver >> 16, ver & 0xFFFF))
        self.f = open(self.path, "r+b")
        self._read_header()

    # ---- reading (gsd_find_chunk / gsd_read_chunk) ----
    def find_chunk(self, frame, name):
        if name not in self.names:
            return None
        k = self._lookup.get((int(frame), self.names.index(name)))
        return None if k is None else self.index[k]

    def read_entry(self, e):
        dt = np.dtype(TYPES[int(e["type"])]).newbyteorder("<")
        n, m = int(e["N"]), int(e["M"])
        size = n * m * dt.itemsize
        if int(e["location"]) == 0 or int(e["location"]) + size > self.file_size:
            raise RuntimeError("Invalid GSD file %s" % self.path)
        self.f.seek(int(e["location"]))
        raw = self.f.read(size)
        if len(raw) != size:
            raise RuntimeError("Error reading GSD file %s" % self.path)
        return np.frombuffer(raw, dt).reshape(n, m).astype(dt.newbyteorder("="))

    def read_chunk(self, frame, name):
        e = self.find_chunk(frame, name)
        if e is None:
            raise KeyError("chunk %s not found in frame %d" % (name, frame))
        return self.read_entry(e)


# ---- HOOMD schema ------------------------------------------------------------------------------------------------------------
CATEGORIES = ("attribute", "property", "momentum", "topology")
# chunk -> (dtype, M, default) of the per-particle chunks that are only written when some value is not the default
# (GSDDumpWriter::writeAttributes / writeProperties / writeMomenta)
PARTICLE_CHUNKS = {
    "attribute": [("particles/typeid", np.uint32, 1, 0), ("particles/mass", np.float32, 1, 1.0),
                  ("particles/charge", np.float32, 1, 0.0), ("particles/diameter", np.float32, 1, 1.0),
                  ("particles/body", np.int32, 1, -1), ("particles/moment_inertia", np.float32, 3, 0.0)],
    "property": [("particles/position", np.float32, 3, None), ("particles/orientation", np.float32, 4, (1, 0, 0, 0))],
    "momentum": [("particles/velocity", np.float32, 3, 0.0), ("particles/angmom", np.float32, 4, 0.0),
                 ("particles/image", np.int32, 3, 0)],
}


def _type_block(names):
    """GSDDumpWriter::writeTypeMapping: NUL-padded rows of max_len + 1 bytes"""
    width = max([len(n.encode()) for n in names] + [0]) + 1
    out = np.zeros((len(names), width), np.uint8)
    for k, n in enumerate(names):
        b = n.encode()
        out[k, :len(b)] = np.frombuffer(b, np.uint8)
    return out


def _type_names(block):
    return [bytes(row).split(b"\0")[0].decode() for row in block]


class HoomdGSDWriter:
    """GSDDumpWriter: frames of the HOOMD schema appended to `path` (created when missing or `overwrite`)."""

    def __init__(self, path, overwrite=False, truncate=False, dynamic=("property",), application="hoomd_b200"):
        self.path, self.truncate = path, bool(truncate)
        self.dynamic = set(dynamic) | {"property"}
        if overwrite or not os.path.exists(path):
            create(path, application, "hoomd", (1, 2))
        self.file = GSDFile(path, "a")
        if self.file.schema != "hoomd":
            raise RuntimeError("dump.gsd: Invalid schema in " + path)
        if self.file.schema_version >= make_version(2, 0):
            raise RuntimeError("dump.gsd: Invalid schema version in " + path)
        # GSDDumpWriter::populateNonDefault: which optional chunks frame 0 of an existing file holds
        self.nondefault = {}
        if self.file.nframes:
            for cat in PARTICLE_CHUNKS.values():
                for name, _dt, _m, _d in cat:
                    self.nondefault[name] = self.file.find_chunk(0, name) is not None

    def write_frame(self, step, box, frame, dimensions=3):
        """frame: dict with position (N,3) and optionally velocity, image, typeid, mass, charge, diameter, body,
        moment_inertia, orientation, angmom (per particle, in TAG order), types (names), bonds_group (Nb,2),
        bonds_typeid, bond_types.  box = (Lx, Ly, Lz[, xy, xz, yz])."""
        f = self.file
        if self.truncate:
            f.truncate()
        nframes = f.nframes
        N = len(frame["position"])
        # writeFrameHeader
        f.write_chunk("configuration/step", np.array([step], np.uint64))
        if nframes == 0:
            f.write_chunk("configuration/dimensions", np.array([dimensions], np.uint8))
        b = np.zeros(6, np.float32)
        b[:len(box)] = box
        f.write_chunk("configuration/box", b)
        f.write_chunk("particles/N", np.array([N], np.uint32))
        for cat in ("attribute", "property", "momentum"):
            if not (cat in self.dynamic or nframes == 0):
                continue
            if cat == "attribute":
                f.write_chunk("particles/types", _type_block(list(frame.get("types", ["A"]))))
            for name, dt, m, default in PARTICLE_CHUNKS[cat]:
                key = name.split("/")[1]
                if default is None:                      # position: always
                    f.write_chunk(name, np.asarray(frame[key], dt).reshape(N, m))
                    continue
                if key in frame and frame[key] is not None:
                    a = np.asarray(frame[key], dt).reshape(N, m)
                    all_default = bool((a == np.asarray(default, dt)).all())
                else:
                    a, all_default = None, True
                if not all_default or (nframes > 0 and self.nondefault.get(name, False)):
                    if a is None:
                        a = np.broadcast_to(np.asarray(default, dt), (N, m)).copy()
                    f.write_chunk(name, a)
                    if nframes == 0:
                        self.nondefault[name] = True
        if "topology" in self.dynamic or nframes == 0:
            group = frame.get("bonds_group")
            nb = 0 if group is None else len(group)
            if nb > 0:                                   # GSDDumpWriter::writeTopology writes a section only when N > 0
                f.write_chunk("bonds/N", np.array([nb], np.uint32))
                f.write_chunk("bonds/types", _type_block(list(frame.get("bond_types", []))))
                f.write_chunk("bonds/typeid", np.asarray(frame["bonds_typeid"], np.uint32).reshape(nb, 1))
                f.write_chunk("bonds/group", np.asarray(group, np.uint32).reshape(nb, 2))
        f.end_frame()

    def close(self):
        self.file.close()


def read_hoomd_frame(path, frame=0, from_end=False):
    """GSDReader: one frame as a dict; chunks missing in `frame` are taken from frame 0, else the schema default
    (hoomd/GSDReader.cc:129-371).  from_end: frame counts back from the end (init.read_gsd's restart semantics use 0)."""
    with GSDFile(path, "r") as f:
        if f.schema != "hoomd":
            raise RuntimeError("data.gsd_snapshot: Invalid schema in " + path)
        if f.schema_version >= make_version(2, 0):
            raise RuntimeError("data.gsd_snapshot: Invalid schema version in " + path)
        nframes = f.nframes
        if from_end and frame <= nframes:
            frame = nframes - frame
        if frame >= nframes:
            raise RuntimeError("data.gsd_snapshot: Cannot read frame %d %s only has %d frames" % (frame, path, nframes))

        def chunk(name, cur_n=0):
            e = f.find_chunk(frame, name)
            if e is None and frame != 0:
                e = f.find_chunk(0, name)
            if e is None or (cur_n != 0 and int(e["N"]) != cur_n):
                return None
            return f.read_entry(e)

        out = {}
        c = chunk("configuration/step")
        out["step"] = int(c[0, 0]) if c is not None else 0
        c = chunk("configuration/dimensions")
        out["dimensions"] = int(c[0, 0]) if c is not None else 3
        c = chunk("configuration/box")
        out["box"] = c[:, 0].astype(np.float32) if c is not None else np.array([1, 1, 1, 0, 0, 0], np.float32)
        c = chunk("particles/N")
        N = int(c[0, 0]) if c is not None else 0
        if N == 0:
            raise RuntimeError("data.gsd_snapshot: cannot read a file with 0 particles")
        out["N"] = N
        c = chunk("particles/types")
        out["types"] = _type_names(c) if c is not None else ["A"]
        for cat in PARTICLE_CHUNKS.values():
            for name, dt, m, default in cat:
                c = chunk(name, N)
                key = name.split("/")[1]
                if c is None:
                    c = np.broadcast_to(np.asarray(0 if default is None else default, dt), (N, m)).copy()
                elif c.shape != (N, m) or c.dtype != np.dtype(dt):
                    raise RuntimeError("data.gsd_snapshot: unexpected shape / type of " + name)
                out[key] = c[:, 0] if m == 1 else c
        c = chunk("bonds/N")
        nb = int(c[0, 0]) if c is not None else 0
        out["bonds_N"] = nb
        if nb:
            out["bond_types"] = _type_names(chunk("bonds/types"))
            out["bonds_typeid"] = chunk("bonds/typeid", nb)[:, 0]
            out["bonds_group"] = chunk("bonds/group", nb)
        else:
            out["bond_types"], out["bonds_typeid"] = [], np.zeros(0, np.uint32)
            out["bonds_group"] = np.zeros((0, 2), np.uint32)
        return out
