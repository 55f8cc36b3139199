"""Python binding of the self-contained HDF5 writer/reader (``csrc/runtime/hdf5.cc``).

``File`` offers the small slice of the h5py API that Keras' weight (de)serialisation uses:
groups, contiguous numeric datasets, and attributes that are numbers, byte strings or arrays of
fixed-length byte strings.  h5py / libhdf5 are not installed on this box.
"""
from __future__ import annotations

import ctypes

import numpy as np

from ..ops import _lib

_DT_TO_NP = {0: np.float32, 1: np.float64, 2: np.int32, 3: np.int64, 4: np.uint8}
_NP_TO_DT = {np.dtype(np.float32): 0, np.dtype(np.float64): 1, np.dtype(np.int32): 2,
             np.dtype(np.int64): 3, np.dtype(np.uint8): 4}
_STR = 5


def _rt():
    lib = _lib.runtime()
    if not getattr(lib, "_h5_declared", False):
        vp, cp, i, sz = ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_size_t
        u64p = ctypes.POINTER(ctypes.c_uint64)
        ip = ctypes.POINTER(ctypes.c_int)
        szp = ctypes.POINTER(ctypes.c_size_t)
        lib.cb2rt_last_error.restype = cp
        lib.cb2h5_writer_new.restype = vp
        lib.cb2h5_writer_free.argtypes = [vp]
        lib.cb2h5_writer_add_group.argtypes = [vp, cp]
        lib.cb2h5_writer_add_dataset.argtypes = [vp, cp, i, i, u64p, vp]
        lib.cb2h5_writer_add_attr.argtypes = [vp, cp, cp, i, sz, i, u64p, vp]
        lib.cb2h5_writer_save.argtypes = [vp, cp]
        lib.cb2h5_reader_open.restype = vp
        lib.cb2h5_reader_open.argtypes = [cp]
        lib.cb2h5_reader_close.argtypes = [vp]
        lib.cb2h5_reader_list.argtypes = [vp, cp, cp, sz]
        lib.cb2h5_reader_dataset_info.argtypes = [vp, cp, ip, ip, u64p]
        lib.cb2h5_reader_read_dataset.argtypes = [vp, cp, vp, sz]
        lib.cb2h5_reader_attr_names.argtypes = [vp, cp, cp, sz]
        lib.cb2h5_reader_attr_info.argtypes = [vp, cp, cp, ip, szp, ip, u64p, szp]
        lib.cb2h5_reader_attr_read.argtypes = [vp, cp, cp, vp, sz]
        lib._h5_declared = True
    return lib


def _err():
    return _rt().cb2rt_last_error().decode(errors="replace")


class AttributeTooLarge(ValueError):
    pass


class _Attrs:
    def __init__(self, f: "File", path: str):
        self._f, self._path = f, path

    def __setitem__(self, name, value):
        self._f._set_attr(self._path, name, value)

    def __getitem__(self, name):
        return self._f._get_attr(self._path, name)

    def __contains__(self, name):
        return name in self.keys()

    def keys(self):
        return self._f._attr_names(self._path)

    def get(self, name, default=None):
        return self[name] if name in self else default


class Group:
    def __init__(self, f: "File", path: str):
        self._f, self._path = f, path.rstrip("/") or "/"
        self.attrs = _Attrs(f, self._path)

    @property
    def name(self):
        return self._path

    def _join(self, name):
        return (self._path.rstrip("/") + "/" + name.strip("/")) if name else self._path

    def create_group(self, name):
        self._f._add_group(self._join(name))
        return Group(self._f, self._join(name))

    def require_group(self, name):
        return self.create_group(name)

    def create_dataset(self, name, data=None, shape=None, dtype=None):
        arr = np.asarray(data) if data is not None else np.zeros(shape, dtype or np.float32)
        if dtype is not None:
            arr = arr.astype(dtype)
        self._f._add_dataset(self._join(name), arr)
        return arr

    def keys(self):
        return [n for _k, n in self._f._list(self._path)]

    def items(self):
        return [(n, self[n]) for n in self.keys()]

    def __contains__(self, name):
        try:
            parts = name.strip("/").split("/")
            path = self._path
            for p in parts:
                if p not in [n for _k, n in self._f._list(path)]:
                    return False
                path = path.rstrip("/") + "/" + p
            return True
        except (KeyError, OSError):
            return False

    def __getitem__(self, name):
        path = self._join(name)
        parent, _, leaf = path.rstrip("/").rpartition("/")
        for kind, n in self._f._list(parent or "/"):
            if n == leaf:
                return Group(self._f, path) if kind == "g" else self._f._read_dataset(path)
        raise KeyError(f"no object {path!r} in file")


class File(Group):
    """``File(path, 'w')`` buffers everything and writes on close; ``File(path, 'r')`` reads."""

    def __init__(self, filename, mode="r"):
        self.filename = str(filename)
        self.mode = mode
        lib = _rt()
        self._w = self._r = None
        if mode == "w":
            self._w = lib.cb2h5_writer_new()
        elif mode == "r":
            self._r = lib.cb2h5_reader_open(self.filename.encode())
            if not self._r:
                raise OSError(f"Unable to open file: {_err()}")
        else:
            raise ValueError("mode must be 'r' or 'w'")
        super().__init__(self, "/")

    # -- writer ---------------------------------------------------------------------------------
    def _add_group(self, path):
        if _rt().cb2h5_writer_add_group(self._w, path.encode()):
            raise OSError(_err())

    def _add_dataset(self, path, arr):
        arr = np.asarray(arr, order="C")        # (np.ascontiguousarray would turn a 0-d array into 1-d)
        if arr.dtype not in _NP_TO_DT:
            arr = arr.astype(np.float32)
        dims = (ctypes.c_uint64 * max(arr.ndim, 1))(*arr.shape)
        if _rt().cb2h5_writer_add_dataset(self._w, path.encode(), _NP_TO_DT[arr.dtype], arr.ndim, dims,
                                          arr.ctypes.data_as(ctypes.c_void_p)):
            raise OSError(_err())

    def _set_attr(self, path, name, value):
        lib = _rt()
        if isinstance(value, str):
            value = value.encode("utf8")
        if isinstance(value, (bytes, np.bytes_)):
            data = bytes(value)
            slen = max(len(data), 1)
            buf = ctypes.create_string_buffer(data, slen)
            rc = lib.cb2h5_writer_add_attr(self._w, path.encode(), name.encode(), _STR, slen, 0, None, buf)
        else:
            arr = np.asarray(value)
            if arr.dtype.kind in ("S", "U", "O"):
                items = [v if isinstance(v, bytes) else str(v).encode("utf8") for v in arr.reshape(-1)]
                slen = max([len(s) for s in items] + [1])
                packed = b"".join(s.ljust(slen, b"\0") for s in items)
                dims = (ctypes.c_uint64 * 1)(len(items))
                buf = ctypes.create_string_buffer(packed, max(len(packed), 1))
                rc = lib.cb2h5_writer_add_attr(self._w, path.encode(), name.encode(), _STR, slen, 1,
                                               dims, buf)
            else:
                if arr.dtype == np.bool_:
                    arr = arr.astype(np.uint8)
                if arr.dtype not in _NP_TO_DT:
                    arr = arr.astype(np.float64 if arr.dtype.kind == "f" else np.int64)
                arr = np.ascontiguousarray(arr)
                dims = (ctypes.c_uint64 * max(arr.ndim, 1))(*arr.shape)
                rc = lib.cb2h5_writer_add_attr(self._w, path.encode(), name.encode(),
                                               _NP_TO_DT[arr.dtype], 0, arr.ndim, dims,
                                               arr.ctypes.data_as(ctypes.c_void_p))
        if rc == 2:
            raise AttributeTooLarge(_err())
        if rc:
            raise OSError(_err())

    # -- reader ---------------------------------------------------------------------------------
    def _list(self, path):
        if self._r is None:
            raise OSError("file is open for writing")
        lib = _rt()
        size = 1 << 16
        while True:
            buf = ctypes.create_string_buffer(size)
            n = lib.cb2h5_reader_list(self._r, path.encode(), buf, size)
            if n >= 0:
                break
            if n == -1:
                raise KeyError(_err())
            size = -n + 16
        return [(line[0], line[2:]) for line in buf.value.decode().split("\n") if line]

    def _read_dataset(self, path):
        lib = _rt()
        dt, rank = ctypes.c_int(), ctypes.c_int()
        dims = (ctypes.c_uint64 * 8)()
        if lib.cb2h5_reader_dataset_info(self._r, path.encode(), ctypes.byref(dt), ctypes.byref(rank), dims):
            raise KeyError(_err())
        if dt.value not in _DT_TO_NP:
            raise TypeError(f"dataset {path} has an unsupported datatype")
        shape = tuple(int(dims[i]) for i in range(rank.value))
        arr = np.empty(shape, dtype=_DT_TO_NP[dt.value])
        if lib.cb2h5_reader_read_dataset(self._r, path.encode(), arr.ctypes.data_as(ctypes.c_void_p),
                                         arr.nbytes):
            raise OSError(_err())
        return arr

    def _attr_names(self, path):
        if self._r is None:
            return []
        lib = _rt()
        buf = ctypes.create_string_buffer(1 << 16)
        n = lib.cb2h5_reader_attr_names(self._r, path.encode(), buf, 1 << 16)
        if n < 0:
            raise KeyError(_err())
        return [s for s in buf.value.decode().split("\n") if s]

    def _get_attr(self, path, name):
        lib = _rt()
        dt, rank = ctypes.c_int(), ctypes.c_int()
        slen, nbytes = ctypes.c_size_t(), ctypes.c_size_t()
        dims = (ctypes.c_uint64 * 8)()
        if lib.cb2h5_reader_attr_info(self._r, path.encode(), name.encode(), ctypes.byref(dt),
                                      ctypes.byref(slen), ctypes.byref(rank), dims, ctypes.byref(nbytes)):
            raise KeyError(_err())
        raw = ctypes.create_string_buffer(max(nbytes.value, 1))
        if lib.cb2h5_reader_attr_read(self._r, path.encode(), name.encode(), raw, nbytes.value):
            raise OSError(_err())
        data = raw.raw[:nbytes.value]
        shape = tuple(int(dims[i]) for i in range(rank.value))
        if dt.value == _STR:
            n = slen.value
            items = [data[i:i + n].rstrip(b"\0") for i in range(0, len(data), n)] if n else []
            if rank.value == 0:
                return items[0] if items else b""
            return np.array(items, dtype=f"S{max(n, 1)}").reshape(shape)
        if dt.value not in _DT_TO_NP:
            raise TypeError(f"attribute {name} has an unsupported datatype")
        arr = np.frombuffer(data, dtype=_DT_TO_NP[dt.value]).reshape(shape)
        return arr[()] if rank.value == 0 else arr.copy()

    # -- lifecycle ------------------------------------------------------------------------------
    def flush(self):
        pass

    def close(self):
        lib = _rt()
        if self._w is not None:
            rc = lib.cb2h5_writer_save(self._w, self.filename.encode())
            lib.cb2h5_writer_free(self._w)
            self._w = None
            if rc:
                raise OSError(_err())
        if self._r is not None:
            lib.cb2h5_reader_close(self._r)
            self._r = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        if exc[0] is None:
            self.close()
        else:
            if self._w is not None:
                _rt().cb2h5_writer_free(self._w)
                self._w = None
            if self._r is not None:
                _rt().cb2h5_reader_close(self._r)
                self._r = None

    def __del__(self):
        try:
            if self._r is not None:
                _rt().cb2h5_reader_close(self._r)
                self._r = None
        except Exception:
            pass
