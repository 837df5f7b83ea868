"""tools/h5mini.hpp (the C++ HDF5 subset the drop-in tools use) against the committed reference
model file and against an independent reader (oracle/h5lite.py); CPU only."""
import os
import struct
import subprocess
import sys
import zlib

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
from oracle.h5lite import H5File  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden", "liberty-0.035-0.250-pr.h5")
UNDEF = 0xFFFFFFFFFFFFFFFF


@pytest.fixture(scope="module")
def h5tool(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("h5") / "h5tool")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-Wall", "-o", exe,
                           os.path.join(ROOT, "tools", "h5tool.cpp"), "-lz", "-lpthread"])
    return exe


def raw(h5tool, path, name, dtype, tmp, lo=None, hi=None):
    out = os.path.join(tmp, "raw.bin")
    cmd = [h5tool, "raw", path, name, out] + ([str(lo), str(hi)] if lo is not None else [])
    subprocess.check_call(cmd)
    return np.fromfile(out, dtype)


def test_reads_reference_model_file(h5tool, tmp_path):
    """`w` as pr-learn wrote it (unlimited rows, chunk 1 x 5120, gzip 9: src/pr-learn.cpp:397-407)"""
    ls = subprocess.check_output([h5tool, "ls", GOLDEN]).decode().split()
    assert ls == ["w", "f32", "9", "5120", "chunk", "1", "5120", "deflate", "unlimited"]
    want = H5File(GOLDEN).read("w")
    got = raw(h5tool, GOLDEN, "w", np.float32, str(tmp_path)).reshape(9, 5120)
    assert np.array_equal(got, want)
    part = raw(h5tool, GOLDEN, "w", np.float32, str(tmp_path), 3, 8).reshape(5, 5120)
    assert np.array_equal(part, want[3:8])


@pytest.mark.parametrize("chunk_rows,level", [(0, 0), (1, 9), (4, 0), (128, 1)])
def test_copy_round_trip(h5tool, tmp_path, chunk_rows, level):
    out = str(tmp_path / "copy.h5")
    subprocess.check_call([h5tool, "copy", GOLDEN, out, str(chunk_rows), str(level), "3"])
    a, b = H5File(GOLDEN), H5File(out)
    assert np.array_equal(a.read("w"), b.read("w"))
    info = b.info("w")
    if chunk_rows:
        assert info["chunk"] == (chunk_rows, 5120) and info["maxshape"][0] == UNDEF
        assert bool(info["filters"]) == (level > 0)
    # and our own reader reads what our writer wrote
    got = raw(h5tool, out, "w", np.float32, str(tmp_path)).reshape(9, 5120)
    assert np.array_equal(got, a.read("w"))


@pytest.mark.parametrize("rows,chunk_rows,level", [(1000, 1, 1), (5000, 128, 0), (77, 128, 9), (4160, 1, 0)])
def test_pipeline_shapes_and_deep_btrees(h5tool, tmp_path, rows, chunk_rows, level):
    """every dataset shape of the pipeline; 1000+ chunks need a 2- or 3-level chunk B-tree"""
    out = str(tmp_path / "gen.h5")
    subprocess.check_call([h5tool, "gen", out, str(rows), str(chunk_rows), str(level)])
    f = H5File(out)
    assert sorted(f.keys()) == ["Distance", "Indices", "Label", "Patches", "RingParams"]
    r = np.arange(rows, dtype=np.float32)[:, None] * np.float32(0.5) + np.arange(40, dtype=np.float32)[None, :] * np.float32(0.001)
    assert np.array_equal(f.read("Distance"), r)
    assert np.array_equal(f.read("Label").ravel(), (np.arange(rows) % 3 == 0).astype(np.uint8))
    k = np.arange(rows)
    assert np.array_equal(f.read("Indices"), np.stack([k, k // 2, k + 7, k // 3], 1).astype(np.int32))
    p = np.arange(rows // 8)[:, None]
    i = np.arange(4096)[None, :]
    assert np.array_equal(f.read("Patches").reshape(rows // 8, 4096), ((p * 31 + i * 7 + (i >> 6)) & 255).astype(np.uint8))
    assert np.array_equal(f.read("RingParams").ravel(), 0.125 * np.arange(21) - 1.0 / 3.0)
    assert f.info("Distance")["dtype"] == np.float32 and f.info("RingParams")["dtype"] == np.float64
    # row windows through the C++ reader (crossing chunk and B-tree node boundaries)
    for lo, hi in [(0, rows), (rows // 3, rows // 3 + 1), (rows // 2 - 5, min(rows, rows // 2 + 140))]:
        got = raw(h5tool, out, "Distance", np.float32, str(tmp_path), lo, hi).reshape(hi - lo, 40)
        assert np.array_equal(got, r[lo:hi])
    got = raw(h5tool, out, "Patches", np.uint8, str(tmp_path), 2, 5).reshape(3, 4096)
    assert np.array_equal(got, ((p * 31 + i * 7 + (i >> 6)) & 255).astype(np.uint8)[2:5])


def test_append_and_flush_per_row(h5tool, tmp_path):
    out = str(tmp_path / "w.h5")
    subprocess.check_call([h5tool, "append-test", out])
    f = H5File(out)
    w = f.read("w")
    assert w.shape == (5, 5120) and f.info("w")["maxshape"][0] == UNDEF and f.info("w")["chunk"] == (1, 5120)
    for k in range(5):
        want = np.where(np.arange(5120) % 97 == k, np.float32((k + 1) * 0.25), np.float32(0))
        assert np.array_equal(w[k], want)


def _colchunk_file(path, data, crow, deflate, leaf_entries=3):
    """A file laid out like the reference's comp-uprjdists output: `Distance` f32 N x D with
    {crow, 1} COLUMN chunks (src/comp-uprjdists.cpp:289-290), gzip, multi-level chunk B-tree."""
    n, d = data.shape
    buf = bytearray(4096)
    chunks = []
    for r0 in range(0, n, crow):
        for c in range(d):
            blk = np.zeros((crow, 1), np.float32)
            part = data[r0:r0 + crow, c:c + 1]
            blk[:part.shape[0]] = part
            raw = blk.tobytes()
            z = zlib.compress(raw, 6) if deflate else raw
            chunks.append((len(buf), len(z), r0, c))
            buf += z
            buf += b"\0" * (-len(buf) % 8)

    def key(sz, r0, c):
        return struct.pack("<IIQQQ", sz, 0, r0, c, 0)

    level, depth = [(a, sz, r0, c) for a, sz, r0, c in chunks], 0
    while True:
        nodes = [level[i:i + leaf_entries] for i in range(0, len(level), leaf_entries)]
        up = []
        for k, ent in enumerate(nodes):
            addr = len(buf)
            node = b"TREE" + struct.pack("<BBHQQ", 1, depth, len(ent), UNDEF, UNDEF)
            for a, sz, r0, c in ent:
                node += key(sz, r0, c) + struct.pack("<Q", a)
            nxt = nodes[k + 1][0] if k + 1 < len(nodes) else None
            node += key(0, nxt[2], nxt[3]) if nxt else key(0, chunks[-1][2] + crow, d)
            buf += node
            up.append((addr, ent[0][1], ent[0][2], ent[0][3]))
        if len(nodes) == 1:
            root = up[0][0]
            break
        level, depth = up, depth + 1

    def msg(t, body, flags=0):
        body += b"\0" * (-len(body) % 8)
        return struct.pack("<HHB3x", t, len(body), flags) + body

    msgs = [msg(1, struct.pack("<BBB5xQQQQ", 1, 2, 1, n, d, n, d)),
            msg(3, bytes([0x11, 0x20, 0x1f, 0, 4, 0, 0, 0, 0, 0, 32, 0, 23, 8, 0, 23, 127, 0, 0, 0]), 1),
            msg(5, bytes([2, 3, 2, 1, 0, 0, 0, 0]))]
    if deflate:
        msgs.append(msg(0xb, bytes([1, 1, 0, 0, 0, 0, 0, 0, 1, 0, 8, 0, 1, 0, 1, 0]) + b"deflate\0" + struct.pack("<II", 6, 0)))
    msgs.append(msg(8, struct.pack("<BBBQIII", 3, 2, 3, root, crow, 1, 4)))
    body = b"".join(msgs)
    hdr = struct.pack("<BBHII4x", 1, 0, len(msgs), 1, len(body)) + body
    ROOT_, BTREE, HEAP, HDATA, OHDR = 96, 136, 680, 712, 800
    snod = OHDR + len(hdr) + (-len(hdr) % 8)
    buf[0:8] = b"\x89HDF\r\n\x1a\n"
    buf[13], buf[14] = 8, 8
    struct.pack_into("<HH", buf, 16, 4, 16)
    struct.pack_into("<QQQQ", buf, 24, 0, UNDEF, len(buf), UNDEF)
    struct.pack_into("<QQI4xQQ", buf, 56, 0, ROOT_, 1, BTREE, HEAP)
    struct.pack_into("<BBHII4xHHB3xQQ", buf, ROOT_, 1, 0, 1, 1, 24, 0x11, 16, 0, BTREE, HEAP)
    buf[BTREE:BTREE + 4] = b"TREE"
    struct.pack_into("<BBHQQQQQ", buf, BTREE + 4, 0, 0, 1, UNDEF, UNDEF, 0, snod, 8)
    buf[HEAP:HEAP + 4] = b"HEAP"
    struct.pack_into("<QQQ", buf, HEAP + 8, 88, 24, HDATA)
    buf[HDATA + 8:HDATA + 16] = b"Distance"
    struct.pack_into("<QQ", buf, HDATA + 24, 1, 64)
    buf[OHDR:OHDR + len(hdr)] = hdr
    buf[snod:snod + 4] = b"SNOD"
    struct.pack_into("<BBHQQ", buf, snod + 4, 1, 0, 1, 8, OHDR)
    with open(path, "wb") as f:
        f.write(buf)


@pytest.mark.parametrize("deflate", [False, True])
def test_reads_column_chunked_distance(h5tool, tmp_path, deflate):
    rng = np.random.default_rng(3)
    data = rng.normal(0, 1, (300, 7)).astype(np.float32)
    path = str(tmp_path / "colchunk.h5")
    _colchunk_file(path, data, 128, deflate)
    assert np.array_equal(H5File(path).read("Distance"), data)          # the fixture itself is well-formed
    got = raw(h5tool, path, "Distance", np.float32, str(tmp_path)).reshape(300, 7)
    assert np.array_equal(got, data)
    for lo, hi in [(0, 1), (127, 129), (250, 300), (128, 256)]:
        got = raw(h5tool, path, "Distance", np.float32, str(tmp_path), lo, hi).reshape(hi - lo, 7)
        assert np.array_equal(got, data[lo:hi])


def test_missing_dataset_and_bad_file_fail_loudly(h5tool, tmp_path):
    bad = tmp_path / "bad.h5"
    bad.write_bytes(b"not hdf5" * 100)
    assert subprocess.call([h5tool, "ls", str(bad)], stderr=subprocess.DEVNULL) == 2
    assert subprocess.call([h5tool, "raw", GOLDEN, "nope", str(tmp_path / "x.bin")], stderr=subprocess.DEVNULL) == 2


# ---- structure of what h5mini writes against what libhdf5 wrote -----------------------------------------------------
def _structure(path, name):
    """Low-level facts of one dataset, byte-parsed independently of both readers: superblock, object header,
    the messages with their raw bytes (addresses masked), the chunk B-tree's node type / key layout and the zlib
    framing of every chunk."""
    f = H5File(path)
    b = f.b
    out = {"superblock": (b[8], b[13], b[14]),                          # version, offset size, length size
           "group_leaf_k": struct.unpack_from("<H", b, 16)[0], "group_internal_k": struct.unpack_from("<H", b, 18)[0],
           "root_cache_type": struct.unpack_from("<I", b, 56 + 16)[0]}
    addr = f.datasets[name]
    ver, _, nmsg, refc, hsize = struct.unpack_from("<BBHII", b, addr)
    out["ohdr"] = (ver, refc)
    msgs = {}
    for mtype, data in f._messages(addr):
        if mtype in (0x00, 0x10):                                       # NIL padding / continuation: layout detail
            continue
        msgs[mtype] = bytes(data)
    out["msg_types"] = sorted(msgs)
    sp = msgs[0x01]
    out["dataspace"] = (sp[0], sp[1], sp[2], struct.unpack_from("<%dQ" % (2 * sp[1]), sp, 8))
    out["datatype"] = msgs[0x03][:8 + 12]
    lay = msgs[0x08]
    nd = lay[2]
    out["layout"] = (lay[0], lay[1], nd, struct.unpack_from("<%dI" % nd, lay, 11))
    out["filter"] = msgs.get(0x0B)
    out["fill"] = msgs.get(0x05)
    btree = struct.unpack_from("<Q", lay, 3)[0]
    assert b[btree:btree + 4] == b"TREE"
    ntype, level, used = struct.unpack_from("<BBH", b, btree + 4)
    left, right = struct.unpack_from("<QQ", b, btree + 8)
    out["btree"] = (ntype, level, used, left == UNDEF, right == UNDEF)
    ksz = 8 + 8 * nd
    keys, zhdr = [], set()
    p = btree + 24
    for i in range(used):
        csize, fmask = struct.unpack_from("<II", b, p)
        offs = struct.unpack_from("<%dQ" % nd, b, p + 8)
        child = struct.unpack_from("<Q", b, p + ksz)[0]
        keys.append((fmask, offs))
        zhdr.add(bytes(b[child:child + 2]))
        assert len(zlib.decompress(b[child:child + csize])) == 5120 * 4
        p += ksz + 8
    out["keys"] = keys
    out["final_key"] = struct.unpack_from("<%dQ" % nd, b, p + 8)          # the key behind the last child
    out["zlib_header"] = sorted(zhdr)
    return out


def test_writer_structure_matches_libhdf5_written_file(h5tool, tmp_path):
    """The reference's own `w` file was written by libhdf5 (through OpenCV's hdf module).  h5mini re-writes the same
    dataset (unlimited rows, chunk 1 x 5120, deflate 9); every structural element a libhdf5 reader walks must be the
    same kind of object with the same field values -- only addresses, sizes of the compressed chunks, the
    modification time and the allocation order of the blocks may differ.  No libhdf5 exists in this image, so this
    is the strongest available check that libhdf5 can read h5mini's files (INTEGRATION.md says so)."""
    out = str(tmp_path / "rewritten.h5")
    subprocess.check_call([h5tool, "copy", GOLDEN, out, "1", "9", "2"])
    ref, got = _structure(GOLDEN, "w"), _structure(out, "w")
    assert got["superblock"] == ref["superblock"] == (0, 8, 8)
    assert got["ohdr"] == ref["ohdr"] == (1, 1)
    assert got["root_cache_type"] == ref["root_cache_type"] == 1            # symbol table cached in the root entry
    assert got["dataspace"] == ref["dataspace"]                              # v1, rank 2, max-dims flag, dims, unlimited
    assert got["datatype"] == ref["datatype"]                                # IEEE f32 LE: class, bit fields, properties
    assert got["layout"] == ref["layout"] == (3, 2, 3, (1, 5120, 4))         # v3, chunked, chunk dims + element size
    assert got["filter"] == ref["filter"]                                    # pipeline v1: deflate, client value 9
    assert got["btree"][:2] == ref["btree"][:2] == (1, 0)                    # chunk B-tree, leaf
    assert got["btree"][2:] == ref["btree"][2:]                              # entries used, no siblings
    assert got["keys"] == ref["keys"] and got["final_key"] == ref["final_key"]
    assert got["zlib_header"] == ref["zlib_header"]                          # same zlib framing (78 da)
    # messages libhdf5 wrote that h5mini must write too (a reader may require them); extras are reported
    assert set(ref["msg_types"]) <= set(got["msg_types"]) | {0x12}, (ref["msg_types"], got["msg_types"])
    assert {0x01, 0x03, 0x08, 0x0B} <= set(got["msg_types"])
    if ref["fill"] is not None and got["fill"] is not None:
        assert got["fill"][:4] == ref["fill"][:4]                            # fill value message version / times / defined


def test_reads_every_file_of_the_reference_workspace(h5tool, tmp_path):
    """SURVEY 7 step 1 / VERDICT r1 weak 13: every .h5 file libhdf5 wrote under the reference's workspace (1 181 models:
    pr-learn's appendable, chunked, deflated `w`, pj-learn's contiguous W / A, the rank files) parses with h5mini, with
    the shapes the tools rely on; every tenth one is read in full.  Only where the reference is present (this
    container); the GPU box does not have it."""
    ws = "/root/reference/workspace"
    if not os.path.isdir(ws):
        pytest.skip("reference workspace not present")
    files = sorted(os.path.join(d, f) for d, _, fs in os.walk(ws) for f in fs if f.endswith(".h5"))
    assert len(files) >= 1000
    n_pr = n_pj = 0
    for k, path in enumerate(files):
        out = subprocess.run([h5tool, "ls", path], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert out.returncode == 0, (path, out.stderr.decode())
        sets = {ln.split()[0]: ln.split() for ln in out.stdout.decode().splitlines() if ln.strip()}
        if path.endswith("-pr.h5"):
            n_pr += 1
            assert sets["w"][1] == "f32" and int(sets["w"][3]) == 5120 and "chunk" in sets["w"], path
            name, cols = "w", 5120
        elif path.endswith("-pj.h5"):
            n_pj += 1
            d = int(sets["A"][2])
            assert sets["A"][1:4] == ["f32", str(d), str(d)] and int(sets["W"][3]) == d and 0 < int(sets["W"][2]) <= d, path
            name, cols = "W", d
        else:
            assert sets, path
            continue
        if k % 10 == 0:
            got = raw(h5tool, path, name, np.float32, str(tmp_path))
            assert got.size % cols == 0 and got.size > 0 and np.isfinite(got).all(), path
    assert n_pr >= 700 and n_pj >= 400
