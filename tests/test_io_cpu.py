"""CPU suite: the file formats either side of the path (lidar_odometry_demo_b200/host/lo_io.hpp):
PCD v0.7 reader/writer for the three wire structs, raw 32-byte scan records, TUM / KITTI trajectories. Pure host C++,
compiled here with g++ and driven through tests/csrc/io_test.cpp."""
import os
import struct
import subprocess

import numpy as np
import pytest

from lidar_odometry_demo_b200 import synthetic as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_PCD = "/root/reference/test/test_data/intersection00056.pcd"


@pytest.fixture(scope="module")
def io_test(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("io") / "io_test")
    pkg = os.path.join(ROOT, "lidar_odometry_demo_b200")
    subprocess.run(["g++", "-std=c++20", "-O2", "-Wall", "-o", exe, os.path.join(ROOT, "tests", "csrc", "io_test.cpp"), f"-L{pkg}",
                    "-l:liblo_b200.so", f"-Wl,-rpath,{pkg}"], check=True)
    return exe


def test_round_trips(io_test, tmp_path):
    r = subprocess.run([io_test, "roundtrip", str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 0 and "PASSED" in r.stdout, r.stdout + r.stderr
    # the binary files are what PCL would write for these structs: records unchanged behind the header
    blob = open(tmp_path / "c_bin.pcd", "rb").read()
    head, data = blob.split(b"DATA binary\n", 1)
    assert b"FIELDS x y z _ intensity ring _ time _" in head and len(data) == 1237 * 32
    assert data == open(tmp_path / "c.bin", "rb").read()


def lzf_compress(raw):
    """Small LZF encoder (literal runs + back references), enough to exercise the decoder."""
    out, lit, i, n = bytearray(), bytearray(), 0, len(raw)

    def flush():
        for s in range(0, len(lit), 32):
            chunk = lit[s:s + 32]
            out.append(len(chunk) - 1)
            out.extend(chunk)
        lit.clear()

    last = {}
    while i < n:
        key = bytes(raw[i:i + 3])
        j = last.get(key)
        last[key] = i
        if j is not None and len(key) == 3 and 0 < i - j <= 8192:
            length = 3
            while i + length < n and raw[j + length] == raw[i + length] and length < 264:
                length += 1
            flush()
            off, ln = i - j - 1, length - 2
            if ln < 7:
                out.append((ln << 5) | (off >> 8))
            else:
                out.append((7 << 5) | (off >> 8))
                out.append(ln - 7)
            out.append(off & 0xFF)
            i += length
        else:
            lit.append(raw[i])
            i += 1
    flush()
    return bytes(out)


def test_binary_compressed_and_field_orders(io_test, tmp_path):
    """binary_compressed stores field after field; fields may come in any order, with padding and other types."""
    n = 700
    rng = np.random.default_rng(5)
    x, y, z = (rng.normal(size=n) * 30).astype(np.float32), (rng.normal(size=n) * 30).astype(np.float32), np.zeros(n, np.float32)
    z[::3] = 1.25                                   # repetitive column: back references, overlapping ones included
    ring = (np.arange(n) % 16).astype(np.uint16)
    t = (np.arange(n) / n).astype(np.float64)       # time as F8
    inten = rng.integers(0, 255, n).astype(np.uint8)
    header = ("# .PCD v0.7\nVERSION 0.7\nFIELDS t ring _ z y x intensity\nSIZE 8 2 1 4 4 4 1\nTYPE F U U F F F U\nCOUNT 1 1 3 1 1 1 1\n"
              f"WIDTH {n}\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS {n}\nDATA binary_compressed\n").encode()
    soa = t.tobytes() + ring.tobytes() + bytes(3 * n) + z.tobytes() + y.tobytes() + x.tobytes() + inten.tobytes()
    comp = lzf_compress(soa)
    assert len(comp) < len(soa)
    path = tmp_path / "scan_c.pcd"
    path.write_bytes(header + struct.pack("<II", len(comp), len(soa)) + comp)
    r = subprocess.run([io_test, "xyzirt", str(path), str(tmp_path / "out.bin")], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.strip() == str(n), r.stdout + r.stderr
    got = np.fromfile(tmp_path / "out.bin", S.XYZIRT)
    assert np.array_equal(got["x"], x) and np.array_equal(got["y"], y) and np.array_equal(got["z"], z)
    assert np.array_equal(got["ring"], ring) and np.array_equal(got["time"], t.astype(np.float32))
    assert np.array_equal(got["intensity"], inten.astype(np.float32))
    # the same table as DATA binary (array of records)
    rec = np.zeros(n, np.dtype([("t", "<f8"), ("ring", "<u2"), ("pad", "u1", 3), ("z", "<f4"), ("y", "<f4"), ("x", "<f4"), ("i", "u1")]))
    rec["t"], rec["ring"], rec["z"], rec["y"], rec["x"], rec["i"] = t, ring, z, y, x, inten
    assert rec.dtype.itemsize == 26
    path2 = tmp_path / "scan_b.pcd"
    path2.write_bytes(header.replace(b"binary_compressed", b"binary") + rec.tobytes())
    r = subprocess.run([io_test, "xyzirt", str(path2), str(tmp_path / "out2.bin")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert open(tmp_path / "out2.bin", "rb").read() == open(tmp_path / "out.bin", "rb").read()


def test_malformed_files_are_rejected(io_test, tmp_path):
    good = ("VERSION 0.7\nFIELDS x y z\nSIZE 4 4 4\nTYPE F F F\nCOUNT 1 1 1\nWIDTH 4\nHEIGHT 1\nPOINTS 4\nDATA binary\n").encode()
    cases = {
        "short.pcd": good + bytes(4 * 12 - 1),
        "nodata.pcd": b"VERSION 0.7\nFIELDS x y z\nSIZE 4 4 4\nTYPE F F F\n",
        "sizes.pcd": good.replace(b"SIZE 4 4 4", b"SIZE 4 4"),
        "kind.pcd": good.replace(b"DATA binary", b"DATA zip") + bytes(48),
        "ascii.pcd": good.replace(b"DATA binary", b"DATA ascii") + b"1 2 3\n4 5\n",
        "noxyz.pcd": good.replace(b"FIELDS x y z", b"FIELDS a b c") + bytes(48),
        "lzf.pcd": good.replace(b"DATA binary", b"DATA binary_compressed") + struct.pack("<II", 3, 48) + b"\xe0\x01\x02",
    }
    for name, blob in cases.items():
        (tmp_path / name).write_bytes(blob)
        r = subprocess.run([io_test, "bad", str(tmp_path / name)], capture_output=True, text=True)
        assert r.returncode == 3, (name, r.returncode, r.stdout)
    r = subprocess.run([io_test, "bad", str(tmp_path / "missing.pcd")], capture_output=True, text=True)
    assert r.returncode == 3


def test_trajectory_formats(io_test, tmp_path):
    r = subprocess.run([io_test, "traj", str(tmp_path / "tum.txt"), str(tmp_path / "kitti.txt")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    tum = np.loadtxt(tmp_path / "tum.txt")
    kitti = np.loadtxt(tmp_path / "kitti.txt")
    assert tum.shape == (2, 8) and kitti.shape == (2, 12)
    assert np.allclose(tum[0], [0, 0, 0, 0, 0, 0, 0, 1]) and np.allclose(kitti[0].reshape(3, 4), np.eye(3, 4))
    # timestamp tx ty tz qx qy qz qw; a quarter turn about z
    assert np.allclose(tum[1], [0.1, 1.5, -2.25, 0.125, 0, 0, 0.70710678, 0.70710678], atol=1e-7)
    assert np.allclose(kitti[1].reshape(3, 4), [[0, -1, 0, 1.5], [1, 0, 0, -2.25], [0, 0, 1, 0.125]], atol=1e-6)


@pytest.mark.skipif(not os.path.exists(REF_PCD), reason="the reference tree is only mounted in the build container")
def test_reads_the_reference_fixture(io_test, tmp_path):
    """test/test_data/intersection00056.pcd (PCD v0.7 binary, `rgb _ x y z _`, 32-byte records) against the committed
    extraction tests/golden/pcd_intersection00056.npz (tools/make_golden.py)."""
    r = subprocess.run([io_test, "xyz", REF_PCD, str(tmp_path / "xyz.f32")], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.strip() == "59691", r.stdout + r.stderr
    got = np.fromfile(tmp_path / "xyz.f32", np.float32).reshape(-1, 3)
    want = np.load(os.path.join(ROOT, "tests", "golden", "pcd_intersection00056.npz"))["xyz"].astype(np.float32)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
