"""TensorBoard event files (``events.out.tfevents.*``) without TensorFlow.

The reference's examples log with ``tf.keras.callbacks.TensorBoard`` and its README promises
"stream logs and monitor them on hosted TensorBoard" (README.md:415-416;
core/tests/examples/call_run_within_script_with_keras_fit.py:94-99); cloud_fit's tests check that
the log directory appears (experimental/cloud_fit/tests/unit/remote_test.py:72,96).  So that a real
TensorBoard can open what this framework writes, the callback emits the TFRecord-framed ``Event``
protocol buffers TensorBoard reads, encoded by hand:

    record  = uint64 length | uint32 masked_crc32c(length) | bytes data | uint32 masked_crc32c(data)
    Event   = { 1: double wall_time, 2: int64 step, 3: string file_version | 5: Summary summary }
    Summary = { 1: repeated Value }      Value = { 1: string tag, 2: float simple_value | 5: HistogramProto histo }
    HistogramProto = { 1: min, 2: max, 3: num, 4: sum, 5: sum_squares (doubles),
                       6: packed double bucket_limit[], 7: packed double bucket[] }

``read_events`` is the matching reader (used by the tests and by ``DistributingCloudTuner`` style
log scraping); it verifies both checksums of every record.
"""
from __future__ import annotations

import os
import socket
import struct
import time

import numpy as np

_CRC_TABLE = None


def _table():
    global _CRC_TABLE
    if _CRC_TABLE is None:
        poly = 0x82F63B78                       # CRC-32C (Castagnoli), reflected
        table = []
        for i in range(256):
            c = i
            for _ in range(8):
                c = (c >> 1) ^ poly if c & 1 else c >> 1
            table.append(c)
        _CRC_TABLE = table
    return _CRC_TABLE


def crc32c(data: bytes) -> int:
    table = _table()
    c = 0xFFFFFFFF
    for b in data:
        c = table[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def masked_crc32c(data: bytes) -> int:
    c = crc32c(data)
    return ((((c >> 15) | (c << 17)) & 0xFFFFFFFF) + 0xA282EAD8) & 0xFFFFFFFF


# -- protobuf wire format (the few field kinds needed) -------------------------------------------
def _varint(n: int) -> bytes:
    n &= (1 << 64) - 1
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _key(field: int, wire: int) -> bytes:
    return _varint((field << 3) | wire)


def _f_double(field, v):
    return _key(field, 1) + struct.pack("<d", float(v))


def _f_float(field, v):
    return _key(field, 5) + struct.pack("<f", float(v))


def _f_varint(field, v):
    return _key(field, 0) + _varint(int(v))


def _f_bytes(field, b: bytes):
    return _key(field, 2) + _varint(len(b)) + b


def _histogram(values, bins=30) -> bytes:
    v = np.asarray(values, dtype=np.float64).reshape(-1)
    if v.size == 0:
        v = np.zeros(1)
    counts, edges = np.histogram(v, bins=bins)
    msg = _f_double(1, v.min()) + _f_double(2, v.max()) + _f_double(3, v.size) + _f_double(4, v.sum()) + \
        _f_double(5, float((v * v).sum()))
    msg += _f_bytes(6, struct.pack("<{}d".format(len(counts)), *edges[1:]))      # right edges
    msg += _f_bytes(7, struct.pack("<{}d".format(len(counts)), *counts.astype(np.float64)))
    return msg


def _event(wall_time, step, payload_field=None, payload=b"") -> bytes:
    msg = _f_double(1, wall_time) + _f_varint(2, step)
    if payload_field is not None:
        msg += _f_bytes(payload_field, payload)
    return msg


class EventFileWriter:
    """Appends events to ``<logdir>/events.out.tfevents.<time>.<host>.<pid>.<n>``."""

    _count = 0

    def __init__(self, logdir: str, filename_suffix: str = ""):
        os.makedirs(logdir, exist_ok=True)
        EventFileWriter._count += 1
        name = "events.out.tfevents.{:010d}.{}.{}.{}{}".format(
            int(time.time()), socket.gethostname(), os.getpid(), EventFileWriter._count, filename_suffix)
        self.path = os.path.join(logdir, name)
        self._f = open(self.path, "ab")
        self._record(_event(time.time(), 0, 3, b"brain.Event:2"))

    def _record(self, data: bytes) -> None:
        header = struct.pack("<Q", len(data))
        self._f.write(header + struct.pack("<I", masked_crc32c(header)) + data +
                      struct.pack("<I", masked_crc32c(data)))

    def add_scalar(self, tag: str, value: float, step: int, wall_time: float | None = None) -> None:
        val = _f_bytes(1, tag.encode("utf8")) + _f_float(2, value)
        self._record(_event(wall_time or time.time(), step, 5, _f_bytes(1, val)))

    def add_histogram(self, tag: str, values, step: int, bins: int = 30, wall_time: float | None = None) -> None:
        val = _f_bytes(1, tag.encode("utf8")) + _f_bytes(5, _histogram(values, bins))
        self._record(_event(wall_time or time.time(), step, 5, _f_bytes(1, val)))

    def flush(self) -> None:
        self._f.flush()

    def close(self) -> None:
        if not self._f.closed:
            self._f.close()


# -- reader ------------------------------------------------------------------------------------------
def _parse(msg: bytes) -> dict:
    """field number -> list of raw values (ints for varint / fixed, bytes for length-delimited)."""
    out, i = {}, 0
    while i < len(msg):
        key, shift = 0, 0
        while True:
            b = msg[i]
            i += 1
            key |= (b & 0x7F) << shift
            shift += 7
            if not b & 0x80:
                break
        field, wire = key >> 3, key & 7
        if wire == 0:
            v, shift = 0, 0
            while True:
                b = msg[i]
                i += 1
                v |= (b & 0x7F) << shift
                shift += 7
                if not b & 0x80:
                    break
        elif wire == 1:
            v = msg[i:i + 8]
            i += 8
        elif wire == 5:
            v = msg[i:i + 4]
            i += 4
        elif wire == 2:
            n, shift = 0, 0
            while True:
                b = msg[i]
                i += 1
                n |= (b & 0x7F) << shift
                shift += 7
                if not b & 0x80:
                    break
            v = msg[i:i + n]
            i += n
        else:
            raise ValueError("unsupported wire type {}".format(wire))
        out.setdefault(field, []).append(v)
    return out


def read_records(path: str):
    """Yields the payload of every record; raises ``ValueError`` on a checksum mismatch."""
    with open(path, "rb") as f:
        while True:
            header = f.read(8)
            if not header:
                return
            if len(header) < 8:
                raise ValueError("truncated record header in " + path)
            (length,) = struct.unpack("<Q", header)
            (crc,) = struct.unpack("<I", f.read(4))
            if crc != masked_crc32c(header):
                raise ValueError("corrupted record length in " + path)
            data = f.read(length)
            (crc,) = struct.unpack("<I", f.read(4))
            if len(data) != length or crc != masked_crc32c(data):
                raise ValueError("corrupted record in " + path)
            yield data


def read_events(path: str) -> list:
    """``[{"wall_time", "step", "file_version"? , "scalars": {tag: value}, "histograms": {tag: {...}}}]``."""
    events = []
    for data in read_records(path):
        ev = _parse(data)
        out = {"wall_time": struct.unpack("<d", ev[1][0])[0] if 1 in ev else 0.0,
               "step": ev.get(2, [0])[0], "scalars": {}, "histograms": {}}
        if 3 in ev:
            out["file_version"] = ev[3][0].decode()
        for summary in ev.get(5, []):
            for value in _parse(summary).get(1, []):
                v = _parse(value)
                tag = v[1][0].decode("utf8")
                if 2 in v:
                    out["scalars"][tag] = struct.unpack("<f", v[2][0])[0]
                if 5 in v:
                    h = _parse(v[5][0])
                    d = {name: struct.unpack("<d", h[f][0])[0] for f, name in
                         ((1, "min"), (2, "max"), (3, "num"), (4, "sum"), (5, "sum_squares"))}
                    for f, name in ((6, "bucket_limit"), (7, "bucket")):
                        raw = h.get(f, [b""])[0]
                        d[name] = list(struct.unpack("<{}d".format(len(raw) // 8), raw))
                    out["histograms"][tag] = d
        events.append(out)
    return events


def event_files(logdir: str) -> list:
    return sorted(os.path.join(logdir, f) for f in os.listdir(logdir) if f.startswith("events.out.tfevents."))
