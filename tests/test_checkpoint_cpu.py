"""Checkpoint / frame-dump formats of the reference (src/Callbacks.cpp, src/NbodySystem.cpp:41,56-67): byte layout,
round trip, JSON text, file rolling.  CPU only."""
import os
import struct

import numpy as np


def test_params_layout_and_round_trip(pkg, tmp_path):
    ck = pkg.checkpoint
    n = 1234
    rng = np.random.default_rng(1)
    pos = rng.standard_normal((n, 4)).astype(np.float32)
    vel = rng.standard_normal((n, 4)).astype(np.float32)
    path = str(tmp_path / "run_10.params")
    ck.save_params(path, pos, vel)
    assert os.path.getsize(path) == 28 * n                          # N = file_size / 28
    raw = open(path, "rb").read()
    first = struct.unpack("<7f", raw[:28])                          # x y z m vx vy vz, little endian
    assert first == tuple(float(v) for v in (*pos[0], *vel[0, :3]))
    p2, v2 = ck.load_params(path)
    assert np.array_equal(p2, pos) and np.array_equal(v2[:, :3], vel[:, :3]) and np.all(v2[:, 3] == 0)
    open(path, "ab").write(b"\x00" * 13)                            # a trailing partial record is ignored
    p3, _ = ck.load_params(path)
    assert p3.shape[0] == n


def test_config_json_is_the_reference_text(pkg, tmp_path):
    ck = pkg.checkpoint
    assert ck.config_json_text(32768, True) == '{\n\t"N": "32768",\n\t"Space": "R2\n}\n'
    s = ck.CheckpointSaver(5, "cfg", str(tmp_path))
    s.reset(100, False)
    assert open(tmp_path / "configs" / "cfg.json").read() == '{\n\t"N": "100",\n\t"Space": "R3\n}\n'
    pos = np.ones((100, 4), np.float32)
    vel = np.zeros((100, 4), np.float32)
    for tick in range(10):
        s.end(tick, pos, vel)
    assert sorted(os.listdir(tmp_path / "configs")) == ["cfg.json", "cfg_10.params", "cfg_5.params"]


def test_frame_dumps_roll_over(pkg, tmp_path):
    ck = pkg.checkpoint
    n = 100000
    pos = np.arange(n * 4, dtype=np.float32).reshape(n, 4)
    w = ck.BinaryDataSaver(2, "", 2, str(tmp_path))                 # 2 MB per file: one 1.2 MB frame each
    w.reset(pos)
    for tick in range(6):
        w.end(tick, pos)
    w.close()
    assert np.array_equal(np.fromfile(tmp_path / "data" / "masses.binary", "<f4"), pos[:, 3])
    files = sorted(os.listdir(tmp_path / "data"))
    assert files == ["data0.binary", "data1.binary", "data2.binary", "masses.binary"]
    frame = np.fromfile(tmp_path / "data" / "data1.binary", "<f4").reshape(-1, 3)
    assert frame.shape[0] == n and np.array_equal(frame, pos[:, :3])
