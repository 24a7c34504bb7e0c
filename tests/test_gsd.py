"""GSD container + HOOMD schema (hoomd_b200/gsd_io.py) against the reference's own GSD library (hoomd/extern/gsd.c, compiled
where it lies into oracle/_ref/libgsd_ref.so): files written here are BYTE-IDENTICAL to what gsd.c writes for the same
sequence of chunks (incl. the index moving to the end of the file when it fills up), each side reads the other's files, and
the schema layer (which chunks a frame holds, defaults that are not written, frame-0 fallback, truncate) behaves as
GSDDumpWriter / GSDReader (hoomd/GSDDumpWriter.cc, hoomd/GSDReader.cc).  CPU only."""
import ctypes as C
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "_ref", "libgsd_ref.so")

from hoomd_b200 import gsd_io  # noqa: E402


class _Entry(C.Structure):   # struct gsd_index_entry (hoomd/extern/gsd.h:62-71)
    _fields_ = [("frame", C.c_uint64), ("N", C.c_uint64), ("location", C.c_int64), ("M", C.c_uint32), ("id", C.c_uint16),
                ("type", C.c_uint8), ("flags", C.c_uint8)]


@pytest.fixture(scope="module")
def ref():
    if not os.path.exists(LIB):
        pytest.skip("oracle/_ref/libgsd_ref.so not built")
    L = C.CDLL(LIB)
    L.gsd_create.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_uint32]
    L.gsd_open.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
    L.gsd_write_chunk.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_uint64, C.c_uint32, C.c_uint8, C.c_void_p]
    L.gsd_end_frame.argtypes = [C.c_void_p]
    L.gsd_close.argtypes = [C.c_void_p]
    L.gsd_truncate.argtypes = [C.c_void_p]
    L.gsd_get_nframes.argtypes = [C.c_void_p]
    L.gsd_get_nframes.restype = C.c_uint64
    L.gsd_find_chunk.argtypes = [C.c_void_p, C.c_uint64, C.c_char_p]
    L.gsd_find_chunk.restype = C.POINTER(_Entry)
    L.gsd_read_chunk.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(_Entry)]
    L.gsd_make_version.restype = C.c_uint32
    return L


def _handle():
    return C.create_string_buffer(4096)   # struct gsd_handle is ~400 bytes


def _frames(nframes, seed=0, n0=37):
    """a deterministic sequence of frames: lists of (name, array) with every element type, varying N"""
    rng = np.random.default_rng(seed)
    out = []
    for k in range(nframes):
        n = n0 + (k % 5)
        fr = [("configuration/step", np.array([10 * k], np.uint64)),
              ("configuration/box", rng.uniform(1, 9, 6).astype(np.float32)),
              ("particles/N", np.array([n], np.uint32)),
              ("particles/position", rng.normal(0, 3, (n, 3)).astype(np.float32)),
              ("particles/image", rng.integers(-3, 4, (n, 3)).astype(np.int32))]
        if k % 3 == 0:
            fr.append(("particles/typeid", rng.integers(0, 4, n).astype(np.uint32)))
            fr.append(("log/double", rng.normal(size=(2, 2))))
        if k % 4 == 1:
            fr.append(("particles/types", gsd_io._type_block(["A", "Bee", "C%d" % k])))
            fr.append(("misc/i8", rng.integers(-100, 100, 5).astype(np.int8)))
            fr.append(("misc/u16", rng.integers(0, 60000, (3, 2)).astype(np.uint16)))
            fr.append(("misc/i64", rng.integers(-2**40, 2**40, 2).astype(np.int64)))
            fr.append(("misc/i16", rng.integers(-30000, 30000, 4).astype(np.int16)))
        out.append(fr)
    return out


def _write_ref(L, path, frames, truncate_at=None):
    assert L.gsd_create(path.encode(), b"app under test", b"hoomd", L.gsd_make_version(1, 2)) == 0
    h = _handle()
    assert L.gsd_open(h, path.encode(), 3) == 0     # GSD_OPEN_APPEND, as GSDDumpWriter does
    for k, fr in enumerate(frames):
        if truncate_at is not None and k == truncate_at:
            assert L.gsd_truncate(h) == 0
        for name, a in fr:
            a = np.ascontiguousarray(a)
            n, m = (a.shape[0], 1) if a.ndim == 1 else a.shape
            assert L.gsd_write_chunk(h, name.encode(), gsd_io.TYPE_OF[a.dtype], n, m, 0, a.ctypes.data_as(C.c_void_p)) == 0
        assert L.gsd_end_frame(h) == 0
    assert L.gsd_close(h) == 0


def _write_mine(path, frames, truncate_at=None):
    gsd_io.create(path, "app under test", "hoomd", (1, 2))
    f = gsd_io.GSDFile(path, "a")
    for k, fr in enumerate(frames):
        if truncate_at is not None and k == truncate_at:
            f.truncate()
        for name, a in fr:
            f.write_chunk(name, a)
        f.end_frame()
    f.close()


@pytest.mark.parametrize("nframes,truncate_at", [(1, None), (12, None), (60, None), (9, 6)])
def test_files_byte_identical_to_reference_library(ref, tmp_path, nframes, truncate_at):
    """60 frames x 5-10 chunks fill the 128-entry index several times: the index is doubled and moved to the end of the file"""
    frames = _frames(nframes)
    a, b = str(tmp_path / "ref.gsd"), str(tmp_path / "mine.gsd")
    _write_ref(ref, a, frames, truncate_at)
    _write_mine(b, frames, truncate_at)
    ba, bb = open(a, "rb").read(), open(b, "rb").read()
    assert len(ba) == len(bb)
    assert ba == bb
    if nframes == 60:
        with gsd_io.GSDFile(b) as f:
            assert f.index_allocated > gsd_io.INITIAL_INDEX and f.index_location > 256


def test_each_side_reads_the_other(ref, tmp_path):
    frames = _frames(40, seed=3)
    a, b = str(tmp_path / "ref.gsd"), str(tmp_path / "mine.gsd")
    _write_ref(ref, a, frames)
    _write_mine(b, frames)
    # mine reads the reference's file
    with gsd_io.GSDFile(a) as f:
        assert f.nframes == 40 and f.schema == "hoomd" and f.application == "app under test"
        assert f.schema_version == gsd_io.make_version(1, 2)
        for k, fr in enumerate(frames):
            for name, arr in fr:
                got = f.read_chunk(k, name)
                assert got.dtype == arr.dtype and np.array_equal(got.reshape(arr.shape), arr), (k, name)
            assert f.find_chunk(k, "no/such/chunk") is None
        assert f.find_chunk(40, "particles/N") is None
    # the reference reads mine (read-only open = mmap path of gsd.c)
    h = _handle()
    assert ref.gsd_open(h, b.encode(), 2) == 0
    assert ref.gsd_get_nframes(h) == 40
    for k, fr in enumerate(frames):
        for name, arr in fr:
            e = ref.gsd_find_chunk(h, k, name.encode())
            assert bool(e), (k, name)
            n, m = (arr.shape[0], 1) if arr.ndim == 1 else arr.shape
            assert (e.contents.N, e.contents.M, e.contents.type) == (n, m, gsd_io.TYPE_OF[arr.dtype])
            buf = np.empty(arr.shape, arr.dtype)
            assert ref.gsd_read_chunk(h, buf.ctypes.data_as(C.c_void_p), e) == 0
            assert np.array_equal(buf, arr)
    assert ref.gsd_close(h) == 0


def test_appending_to_an_existing_file(ref, tmp_path):
    """GSDDumpWriter re-opens an existing trajectory in append mode and continues after the last frame"""
    frames = _frames(30, seed=5)
    a, b = str(tmp_path / "ref.gsd"), str(tmp_path / "mine.gsd")
    _write_ref(ref, a, frames[:17])
    _write_mine(b, frames[:17])
    for path, writer in ((a, "ref"), (b, "mine")):
        if writer == "ref":
            h = _handle()
            assert ref.gsd_open(h, path.encode(), 3) == 0
            for fr in frames[17:]:
                for name, arr in fr:
                    arr = np.ascontiguousarray(arr)
                    n, m = (arr.shape[0], 1) if arr.ndim == 1 else arr.shape
                    assert ref.gsd_write_chunk(h, name.encode(), gsd_io.TYPE_OF[arr.dtype], n, m, 0,
                                               arr.ctypes.data_as(C.c_void_p)) == 0
                assert ref.gsd_end_frame(h) == 0
            assert ref.gsd_close(h) == 0
        else:
            f = gsd_io.GSDFile(path, "a")
            assert f.nframes == 17
            for fr in frames[17:]:
                for name, arr in fr:
                    f.write_chunk(name, arr)
                f.end_frame()
            f.close()
    assert open(a, "rb").read() == open(b, "rb").read()


def test_rejects_what_gsd_c_rejects(tmp_path):
    p = str(tmp_path / "junk.gsd")
    open(p, "wb").write(b"\0" * 300)
    with pytest.raises(RuntimeError):
        gsd_io.GSDFile(p)
    open(p, "wb").write(b"short")
    with pytest.raises(RuntimeError):
        gsd_io.GSDFile(p)
    gsd_io.create(p, "x")
    with gsd_io.GSDFile(p) as f:
        assert f.nframes == 0
        with pytest.raises(RuntimeError):
            f.write_chunk("a", np.zeros(3, np.float32))     # read-only
    with gsd_io.GSDFile(p, "a") as f:
        with pytest.raises(ValueError):
            f.write_chunk("a", np.zeros((3, 0), np.float32))  # M == 0
        with pytest.raises(ValueError):
            f.write_chunk("a", np.zeros(3, np.complex64))


def _particles(n, seed, all_default=False):
    rng = np.random.default_rng(seed)
    fr = dict(position=rng.normal(0, 2, (n, 3)).astype(np.float32), types=["A", "B"])
    if not all_default:
        fr.update(velocity=rng.normal(0, 1, (n, 3)).astype(np.float32), image=rng.integers(-2, 3, (n, 3)).astype(np.int32),
                  typeid=rng.integers(0, 2, n).astype(np.uint32), mass=rng.uniform(0.5, 2, n).astype(np.float32),
                  diameter=np.ones(n, np.float32))
    return fr


def test_schema_defaults_dynamic_categories_and_frame0_fallback(ref, tmp_path):
    p = str(tmp_path / "traj.gsd")
    n = 50
    f0 = _particles(n, 1)
    f0.update(bonds_group=np.array([[0, 1], [1, 2], [7, 9]], np.uint32), bonds_typeid=np.array([0, 1, 0], np.uint32),
              bond_types=["polymer", "x"])
    w = gsd_io.HoomdGSDWriter(p, overwrite=True)               # dynamic = property only
    w.write_frame(100, (10, 11, 12), f0)
    f1 = _particles(n, 2)
    w.write_frame(200, (10.5, 11, 12), f1)
    w.close()
    with gsd_io.GSDFile(p) as f:
        names0 = {f.names[int(e["id"])] for e in f.index if e["frame"] == 0}
        names1 = {f.names[int(e["id"])] for e in f.index if e["frame"] == 1}
    # frame 0: everything non-default; diameter == 1 everywhere is NOT written; frame 1: header + properties only
    assert names0 == {"configuration/step", "configuration/dimensions", "configuration/box", "particles/N", "particles/types",
                      "particles/typeid", "particles/mass", "particles/position", "particles/velocity", "particles/image",
                      "bonds/N", "bonds/types", "bonds/typeid", "bonds/group"}
    assert names1 == {"configuration/step", "configuration/box", "particles/N", "particles/position"}
    r0, r1 = gsd_io.read_hoomd_frame(p, 0), gsd_io.read_hoomd_frame(p, 1)
    assert r0["step"] == 100 and r1["step"] == 200 and r1["dimensions"] == 3
    assert np.allclose(r1["box"], [10.5, 11, 12, 0, 0, 0])
    assert np.array_equal(r1["position"], f1["position"]) and np.array_equal(r0["position"], f0["position"])
    # velocity / mass / typeid / bonds of frame 1 fall back to frame 0 (GSDReader::readChunk)
    for k in ("velocity", "mass", "typeid", "image"):
        assert np.array_equal(r1[k], f0[k]), k
    assert np.array_equal(r1["diameter"], np.ones(n, np.float32)) and np.array_equal(r1["charge"], np.zeros(n, np.float32))
    assert r1["types"] == ["A", "B"] and r1["bond_types"] == ["polymer", "x"]
    assert np.array_equal(r1["bonds_group"], f0["bonds_group"]) and r1["bonds_N"] == 3
    with pytest.raises(RuntimeError):
        gsd_io.read_hoomd_frame(p, 2)
    # the reference library sees the same chunks
    h = _handle()
    assert ref.gsd_open(h, p.encode(), 2) == 0
    assert ref.gsd_get_nframes(h) == 2
    e = ref.gsd_find_chunk(h, 0, b"particles/mass")
    buf = np.empty(n, np.float32)
    assert ref.gsd_read_chunk(h, buf.ctypes.data_as(C.c_void_p), e) == 0 and np.array_equal(buf, f0["mass"])
    assert not ref.gsd_find_chunk(h, 1, b"particles/mass") and not ref.gsd_find_chunk(h, 0, b"particles/diameter")
    ref.gsd_close(h)

    # momentum dynamic: velocities of later frames are written (and default ones too once frame 0 had them)
    p2 = str(tmp_path / "mom.gsd")
    w = gsd_io.HoomdGSDWriter(p2, overwrite=True, dynamic=("momentum",))
    w.write_frame(0, (10, 10, 10), f0)
    fz = dict(position=f1["position"], types=["A", "B"], velocity=np.zeros((n, 3), np.float32))
    w.write_frame(5, (10, 10, 10), fz)
    w.close()
    r = gsd_io.read_hoomd_frame(p2, 1)
    assert np.array_equal(r["velocity"], np.zeros((n, 3), np.float32))     # written although all default
    assert np.array_equal(r["mass"], f0["mass"])                            # attribute: still from frame 0

    # all-default frame 0: only the mandatory chunks
    p3 = str(tmp_path / "plain.gsd")
    w = gsd_io.HoomdGSDWriter(p3, overwrite=True)
    w.write_frame(7, (5, 5, 5), _particles(10, 3, all_default=True))
    w.close()
    with gsd_io.GSDFile(p3) as f:
        assert sorted(f.names) == sorted(["configuration/step", "configuration/dimensions", "configuration/box",
                                          "particles/N", "particles/types", "particles/position"])


def test_truncate_keeps_one_frame_and_append_continues(tmp_path):
    p = str(tmp_path / "restart.gsd")
    w = gsd_io.HoomdGSDWriter(p, overwrite=True, truncate=True)
    for k in range(4):
        w.write_frame(1000 * k, (8, 8, 8), _particles(20, 10 + k))
    w.close()
    with gsd_io.GSDFile(p) as f:
        assert f.nframes == 1
    r = gsd_io.read_hoomd_frame(p, 0)
    assert r["step"] == 3000 and np.array_equal(r["velocity"], _particles(20, 13)["velocity"])
    # not overwrite: an existing trajectory is continued
    p2 = str(tmp_path / "traj.gsd")
    w = gsd_io.HoomdGSDWriter(p2, overwrite=True)
    w.write_frame(0, (8, 8, 8), _particles(20, 1))
    w.close()
    w = gsd_io.HoomdGSDWriter(p2, overwrite=False)
    assert w.nondefault["particles/velocity"] and not w.nondefault["particles/charge"]
    w.write_frame(10, (8, 8, 8), _particles(20, 2))
    w.close()
    with gsd_io.GSDFile(p2) as f:
        assert f.nframes == 2


def test_reads_the_reference_tree_fixtures_like_gsd_c(ref):
    """the polymer fixtures of the reference's own Python tests, where the reference tree is present (this container)"""
    d = "/root/reference/hoomd/md/test-py"
    if not os.path.isdir(d):
        pytest.skip("reference tree not present")
    files = sorted(x for x in os.listdir(d) if x.endswith(".gsd"))
    assert files
    for name in files:
        path = os.path.join(d, name)
        h = _handle()
        assert ref.gsd_open(h, path.encode(), 2) == 0
        with gsd_io.GSDFile(path) as f:
            assert f.nframes == ref.gsd_get_nframes(h)
            for e in f.index:
                cname = f.names[int(e["id"])]
                er = ref.gsd_find_chunk(h, int(e["frame"]), cname.encode())
                assert bool(er) and er.contents.location == int(e["location"])
                mine = f.read_entry(e)
                buf = np.empty(mine.shape, mine.dtype)
                assert ref.gsd_read_chunk(h, buf.ctypes.data_as(C.c_void_p), er) == 0
                assert np.array_equal(buf, mine), (name, cname)
        ref.gsd_close(h)
        fr = gsd_io.read_hoomd_frame(path, 0)
        assert fr["N"] == len(fr["position"]) and fr["bonds_N"] == len(fr["bonds_group"])
