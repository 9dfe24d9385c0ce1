"""Checkpoint and frame-dump files in the reference's formats (SURVEY.md sections 5 and 8f-3).

  configs/<name>_<tick>.params   7 little-endian float32 per body: x y z m vx vy vz  (28 bytes)
                                 written by CheckpointSaver::save (src/Callbacks.cpp:122-142), read by the
                                 NbodySystem(string configPath, ...) constructor (src/NbodySystem.cpp:41,56-67)
  configs/<name>.json            N and Space as strings, byte for byte as CheckpointSaver::initConfig emits it
                                 (src/Callbacks.cpp:96-106; tab-indented, last closing quote eaten by its seekp)
  data/masses.binary             N float32 masses                                   (src/Callbacks.cpp:22-28)
  data/data<k>.binary            per saved tick N x (x, y, z) float32, a new file once maxFileMBSize would be passed
                                 (src/Callbacks.cpp:36-62)

Host-side file I/O only; under sharding the caller passes gathered (replicated) positions and all-gathered
velocities and lets rank 0 write.
"""
import os

import numpy as np


def save_params(path, pos_mass, vel):
    pos_mass = np.asarray(pos_mass, np.float32)
    vel = np.asarray(vel, np.float32)
    n = vel.shape[0]
    rec = np.empty((n, 7), "<f4")
    rec[:, 0:4] = pos_mass[:n, 0:4]
    rec[:, 4:7] = vel[:, 0:3]
    rec.tofile(path)


def load_params(path):
    """Returns (pos_mass [N,4], vel [N,4]); N = file_size / 28 as in src/NbodySystem.cpp:41 (a trailing partial
    record is ignored, like the reference's chunked read)."""
    raw = np.fromfile(path, dtype="<f4")
    n = raw.size // 7
    rec = raw[:n * 7].reshape(n, 7)
    pos = np.ascontiguousarray(rec[:, 0:4], dtype=np.float32)
    vel = np.zeros((n, 4), np.float32)
    vel[:, 0:3] = rec[:, 4:7]
    return pos, vel


def config_json_text(n, space_is_r2):
    """Byte for byte what CheckpointSaver::initConfig leaves on disk when built with GCC: it seeks back three
    characters (`",\\n`) and writes `\\n}`, which eats the closing quote of the last value and keeps the trailing
    newline (src/Callbacks.cpp:103-105).  Nothing in the reference ever parses this file."""
    return "{\n\t\"N\": \"%d\",\n\t\"Space\": \"%s\n}\n" % (n, "R2" if space_is_r2 else "R3")


def save_config_json(path, n, space_is_r2):
    with open(path, "w", newline="") as f:
        f.write(config_json_text(n, space_is_r2))


class CheckpointSaver:
    """Callbacks::CheckpointSaver(interval, name): after tick t with (t+1) % interval == 0 writes
    configs/<name>_<t+1>.params (src/Callbacks.cpp:110-120)."""

    def __init__(self, interval, name="config", root="."):
        self.interval, self.name, self.root = interval, name, root

    def reset(self, n, space_is_r2):
        os.makedirs(os.path.join(self.root, "configs"), exist_ok=True)
        save_config_json(os.path.join(self.root, "configs", self.name + ".json"), n, space_is_r2)

    def end(self, tick, pos_mass, vel):
        if (tick + 1) % self.interval == 0:
            save_params(os.path.join(self.root, "configs", "%s_%d.params" % (self.name, tick + 1)), pos_mass, vel)


class BinaryDataSaver:
    """Callbacks::BinaryDataSaver(interval, name, maxFileMBSize): masses once, then xyz frames every `interval`
    ticks, rolling to data<k+1>.binary before a file would exceed the limit (src/Callbacks.cpp:9-62)."""

    def __init__(self, interval, name="", max_file_mb=750, root="."):
        self.interval, self.max_bytes, self.root = interval, max_file_mb * 1024 * 1024, root
        self.file = None

    def reset(self, pos_mass):
        n = pos_mass.shape[0]
        os.makedirs(os.path.join(self.root, "data"), exist_ok=True)
        np.asarray(pos_mass[:, 3], "<f4").tofile(os.path.join(self.root, "data", "masses.binary"))
        self.size_per_write = n * 3 * 4
        self.file_size, self.file_idx = 0, 0
        self._open()

    def _open(self):
        if self.file:
            self.file.close()
        self.file = open(os.path.join(self.root, "data", "data%d.binary" % self.file_idx), "wb")
        self.file_idx += 1

    def end(self, tick, pos_mass):
        if tick % self.interval == 0:
            if self.file_size + self.size_per_write > self.max_bytes:
                self.file_size = 0
                self._open()
            self.file_size += self.size_per_write
            np.ascontiguousarray(np.asarray(pos_mass)[:, 0:3], "<f4").tofile(self.file)

    def close(self):
        if self.file:
            self.file.close()
            self.file = None
