"""Output files of the reference host written through the C-ABI (pure host code: runs without a GPU).

Pins: the byte layout of save_bin_map_file (reference main.cpp:705-735) restated with struct.pack, the loader of the
reference's plot.py (plot.py:26-73, restated below as `load_bin_map_file`), the text save_csv_stat_file prints
(main.cpp:782-825 with the 16-byte number strings of main.cpp:741-757), and the PNG container (zlib / CRC checked by
Python's own zlib).
"""
import os
import struct
import time
import zlib

import numpy as np
import pytest

from bachelors_b200 import solver as S
from bachelors_b200._abi import pf_stats_row
from bachelors_b200.app import save_folder_name
from bachelors_b200.config import read_config

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load_bin_map_file(path):
    """What plot.py:26-73 does: little-endian header, names, then the maps as (ny, nx) float64."""
    data = open(path, "rb").read()
    magic, count, nx, ny = struct.unpack_from("<iiii", data, 0)
    dx, dy, t = struct.unpack_from("<ddd", data, 16)
    (it,) = struct.unpack_from("<q", data, 40)
    pos = 48
    names = []
    for _ in range(count):
        names.append(data[pos:pos + 32].split(b"\0")[0].decode())
        pos += 32
    maps = {}
    for n in names:
        maps[n] = np.frombuffer(data[pos:pos + nx * ny * 8], dtype=np.float64).reshape(ny, nx)
        pos += nx * ny * 8
    assert pos == len(data)
    return dict(magic=magic, nx=nx, ny=ny, dx=dx, dy=dy, time=t, iter=it, maps=maps)


def expected_bytes(nx, ny, dx, dy, it, t, maps):
    out = struct.pack("<iiii", 0x11223344, len(maps), nx, ny) + struct.pack("<ddd", dx, dy, t) + struct.pack("<q", it)
    for n in maps:
        out += n.encode()[:31].ljust(32, b"\0")
    for a in maps.values():
        out += np.ascontiguousarray(a, dtype="<f8").tobytes()
    return out


def test_maps_bin_is_byte_identical_to_the_reference_layout(tmp_path):
    rng = np.random.default_rng(5)
    nx, ny = 37, 11
    maps = {"F": rng.random((ny, nx)), "U": rng.standard_normal((ny, nx)), "a_rather_long_debug_map_name_over_31_chars": rng.random((ny, nx))}
    path = str(tmp_path / "maps_0003.bin")
    S.write_maps_bin(path, 4.0 / nx, 4.0 / ny, 1234567890123, 0.04, maps)
    assert open(path, "rb").read() == expected_bytes(nx, ny, 4.0 / nx, 4.0 / ny, 1234567890123, 0.04, maps)
    got = load_bin_map_file(path)
    assert got["magic"] == 0x11223344 and got["iter"] == 1234567890123 and got["time"] == 0.04
    assert list(got["maps"]) == ["F", "U", "a_rather_long_debug_map_name_over_31_chars"[:31]]
    assert np.array_equal(got["maps"]["U"], maps["U"])


def test_maps_bin_written_in_slabs_equals_the_one_shot_file(tmp_path):
    rng = np.random.default_rng(6)
    nx, ny = 16, 24
    F, U = rng.random((ny, nx)), rng.random((ny, nx))
    one, parts = str(tmp_path / "one.bin"), str(tmp_path / "parts.bin")
    S.write_maps_bin(one, 0.25, 0.5, 7, 1.5, {"F": F, "U": U})
    S.write_maps_bin_header(parts, nx, ny, 0.25, 0.5, 7, 1.5, ["F", "U"])
    for y0, y1 in ((16, 24), (0, 8), (8, 16)):                      # "ranks" finish in any order
        S.write_maps_bin_rows(parts, nx, ny, 2, 1, y0, U[y0:y1])
        S.write_maps_bin_rows(parts, nx, ny, 2, 0, y0, F[y0:y1])
    assert open(one, "rb").read() == open(parts, "rb").read()
    with pytest.raises(S.PfError):
        S.write_maps_bin_rows(parts, nx, ny, 2, 0, 20, F[:8])        # rows past the end


def test_empty_map_list_and_bad_path(tmp_path):
    path = str(tmp_path / "none.bin")
    S.write_maps_bin_header(path, 4, 4, 1.0, 1.0, 0, 0.0, [])
    assert len(open(path, "rb").read()) == 48
    with pytest.raises(S.PfError):
        S.write_maps_bin(str(tmp_path / "no_such_dir" / "x.bin"), 1.0, 1.0, 0, 0.0, {"F": np.zeros((2, 2))})


def _c_float_str(v):
    return ("%f" % np.float32(v))[:15]


def test_stats_csv_text(tmp_path):
    rows = []
    for i in range(3):
        r = pf_stats_row()
        r.time, r.iter, r.Phi_iters, r.T_iters = 5e-6 * i, i, 4 + i, 3
        r.T_delta_L1, r.T_delta_L2, r.T_delta_max, r.T_delta_min = 1.5 * i, 0.25, 1e-3, -2.5e-4
        r.Phi_delta_L1, r.Phi_delta_L2, r.Phi_delta_max, r.Phi_delta_min = 123456789012.0, 2.0, 3.0, -4.0
        r.step_res_L1[0], r.step_res_L2[0], r.step_res_max[0], r.step_res_min[0] = 0.5, 0.25, 0.125, -0.125
        rows.append(r)
    path = str(tmp_path / "stats.csv")
    S.write_stats_csv(path, False, 128, 64, 5e-6, 1, rows[:2])
    S.write_stats_csv(path, True, 128, 64, 5e-6, 1, rows[2:])
    lines = open(path).read().split("\n")
    assert lines[0] == "128,64,0.000005"
    assert lines[1] == ('"time","iter","Phi_iters","T_iters","T_delta_L1","T_delta_L2","T_delta_max","T_delta_min",'
                        '"Phi_delta_L1","Phi_delta_L2","Phi_delta_max","Phi_delta_min",'
                        '"step_res_L1[0]","step_res_L2[0]","step_res_max[0]","step_res_min[0]"')
    assert len(lines) == 6 and lines[5] == ""
    cols = lines[3].split(",")
    assert cols[:4] == ["0.000005", "1", "5", "3"]
    assert cols[4:8] == ["1.500000", "0.250000", "0.001000", "-0.000250"]
    # a 16-byte buffer truncates long numbers (main.cpp:737-748): "%f" of 123456790528.0f is 19 characters
    assert cols[8] == _c_float_str(123456789012.0) and len(cols[8]) == 15
    assert cols[12:] == ["0.500000", "0.250000", "0.125000", "-0.125000"]
    # the reference's loader (plot.py:104-140) needs: first line nx,ny,dt; a header line; numeric cells
    nx, ny, dt = lines[0].split(",")
    assert (int(nx), int(ny), float(dt)) == (128, 64, 5e-6)
    assert all(float(c) == float(c) for c in lines[2].split(","))


def test_png_container_and_pixels(tmp_path):
    rng = np.random.default_rng(7)
    h, w = 70, 300                                                   # 84 070 raw bytes: more than one stored block
    img = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    for flip in (False, True):
        path = str(tmp_path / ("img%d.png" % flip))
        S.write_png(path, img, flip_y=flip)
        data = open(path, "rb").read()
        assert data[:8] == b"\x89PNG\r\n\x1a\n"
        pos, chunks = 8, []
        while pos < len(data):
            (n,) = struct.unpack(">I", data[pos:pos + 4])
            typ, body = data[pos + 4:pos + 8], data[pos + 8:pos + 8 + n]
            (crc,) = struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])
            assert crc == zlib.crc32(typ + body)
            chunks.append((typ, body))
            pos += 12 + n
        assert [c[0] for c in chunks] == [b"IHDR", b"IDAT", b"IEND"]
        assert struct.unpack(">IIBBBBB", chunks[0][1]) == (w, h, 8, 6, 0, 0, 0)
        raw = np.frombuffer(zlib.decompress(chunks[1][1]), dtype=np.uint8).reshape(h, 1 + w * 4)
        assert not raw[:, 0].any()
        px = raw[:, 1:].reshape(h, w, 4)
        assert np.array_equal(px, img[::-1] if flip else img)


def test_save_folder_name_matches_the_reference_format():
    cfg = read_config(os.path.join(ROOT, "benchmark_configs", "config_explicit-rk4_128.ini"))
    t = time.mktime((2026, 11, 4, 9, 8, 7, 0, 0, -1))
    name = save_folder_name(cfg, t)
    # main.cpp:763-770: folder "/" prefix YYYY-MM-DD__HH-MM-SS__solver postfix, month printed 0-based (tm_mon)
    assert name.endswith("2026-10-04__09-08-07__explicit-rk4" + cfg.snapshot_postfix)
    assert name.startswith(cfg.snapshot_folder + "/" + cfg.snapshot_prefix if cfg.snapshot_folder else cfg.snapshot_prefix)
