"""Torch7 binary serialisation (``torch.save`` / ``torch.load`` / ``torch.MemoryFile``), reader and writer.

Why it exists: the reference checkpoints with ``torch.save`` -- the CIFAR loop saves the whole
``nn.Sequential`` every ``saveep`` epochs (/root/reference/asyncsgd/goot.lua:246-254), BiCNN saves
and reloads the flat parameter tensor (/root/reference/BiCNN/bicnn.lua:259-261,590-594) and preloads
``binary_*`` tables (/root/reference/BiCNN/plaunch.lua:218-229); ``mpiT.serialize`` is the same format
in memory (/root/reference/init.lua:111-132).  Files written here load in Torch7 and vice versa.

Grammar (little-endian, ``long`` = 8 bytes), confirmed on the shipped fixtures
``BiCNN/binary_mapWordIdx2WordStr`` / ``binary_mapWordStr2WordIdx``::

    object  := int32 type, payload
    type 0 nil | 1 number: f64 | 2 string: int32 n, n bytes | 5 boolean: int32
    type 3 table : int32 ref; first time: int32 count, count x (object key, object value)
    type 4 torch : int32 ref; first time: string "V 1", string class, class payload
       torch.XTensor  : int32 ndim, int64 size[ndim], int64 stride[ndim], int64 offset(1-based), object storage
       torch.XStorage : int64 n, n raw elements
       other classes  : object (a table of fields)

Lua tables come back as ``dict`` (integral float keys become ``int``); ``T7Object`` carries the class
name of non-tensor torch objects (nn modules).  Python ``list``/``tuple`` are written as 1-based tables.
"""
from __future__ import annotations

import io
import struct
from typing import Any, BinaryIO, Dict

import numpy as np
import torch

TYPE_NIL, TYPE_NUMBER, TYPE_STRING, TYPE_TABLE, TYPE_TORCH, TYPE_BOOLEAN = 0, 1, 2, 3, 4, 5

_STORAGE_DTYPES = {
    "torch.FloatStorage": np.dtype("<f4"), "torch.DoubleStorage": np.dtype("<f8"),
    "torch.LongStorage": np.dtype("<i8"), "torch.IntStorage": np.dtype("<i4"),
    "torch.ShortStorage": np.dtype("<i2"), "torch.ByteStorage": np.dtype("u1"),
    "torch.CharStorage": np.dtype("i1"), "torch.CudaStorage": np.dtype("<f4"),
    "torch.HalfStorage": np.dtype("<f2"),
}
_TENSOR_TO_STORAGE = {k.replace("Storage", "Tensor"): k for k in _STORAGE_DTYPES}
_TORCH_TO_T7 = {torch.float32: "Float", torch.float64: "Double", torch.int64: "Long", torch.int32: "Int",
                torch.int16: "Short", torch.uint8: "Byte", torch.int8: "Char", torch.float16: "Half"}


class T7Object:
    """A torch class instance that is not a tensor/storage (e.g. ``nn.Sequential``)."""

    def __init__(self, classname: str, fields: Any):
        self.classname = classname
        self.fields = fields

    def __repr__(self):
        return f"T7Object({self.classname})"

    def __getitem__(self, k):
        return self.fields[k]


class T7Storage:
    """A bare storage (rare outside tensors); ``array`` is a numpy view of its elements."""

    def __init__(self, classname: str, array: np.ndarray):
        self.classname = classname
        self.array = array


class _Reader:
    def __init__(self, f: BinaryIO):
        self.f = f
        self.memo: Dict[int, Any] = {}

    def _read(self, n):
        b = self.f.read(n)
        if len(b) != n:
            raise EOFError("truncated Torch7 file")
        return b

    def i32(self):
        return struct.unpack("<i", self._read(4))[0]

    def i64(self):
        return struct.unpack("<q", self._read(8))[0]

    def f64(self):
        return struct.unpack("<d", self._read(8))[0]

    def string(self):
        return self._read(self.i32()).decode("latin-1")

    def array(self, n: int, dt: np.dtype) -> np.ndarray:
        return np.frombuffer(self._read(n * dt.itemsize), dtype=dt).copy()

    def obj(self):
        t = self.i32()
        if t == TYPE_NIL:
            return None
        if t == TYPE_NUMBER:
            v = self.f64()
            return int(v) if v.is_integer() and abs(v) < 2 ** 53 else v
        if t == TYPE_STRING:
            return self.string()
        if t == TYPE_BOOLEAN:
            return self.i32() != 0
        if t == TYPE_TABLE:
            ref = self.i32()
            if ref in self.memo:
                return self.memo[ref]
            out: Dict[Any, Any] = {}
            self.memo[ref] = out
            for _ in range(self.i32()):
                k = self.obj()
                out[k] = self.obj()
            return out
        if t == TYPE_TORCH:
            ref = self.i32()
            if ref in self.memo:
                return self.memo[ref]
            version = self.string()
            cls = self.string() if version.startswith("V ") else version   # pre-versioning files
            val = self._torch_payload(cls, ref)
            self.memo[ref] = val
            return val
        raise ValueError(f"unsupported Torch7 type tag {t} (functions are not supported)")

    def _torch_payload(self, cls, ref):
        if cls in _STORAGE_DTYPES:
            n = self.i64()
            return T7Storage(cls, self.array(n, _STORAGE_DTYPES[cls]))
        if cls in _TENSOR_TO_STORAGE:
            nd = self.i32()
            size = [self.i64() for _ in range(nd)]
            stride = [self.i64() for _ in range(nd)]
            offset = self.i64() - 1
            st = self.obj()
            if st is None or nd == 0:
                return torch.empty(0, dtype=torch.from_numpy(np.zeros(1, _STORAGE_DTYPES[_TENSOR_TO_STORAGE[cls]])).dtype)
            base = torch.from_numpy(st.array)
            return torch.as_strided(base, size, stride, offset)
        obj = T7Object(cls, None)
        self.memo[ref] = obj
        obj.fields = self.obj()
        return obj


class _AsciiReader(_Reader):
    """``torch.load(file, 'ascii')`` (THDiskFile ascii mode, used by the CIFAR-10 ``*.t7`` batches that
    /root/reference/asyncsgd/goot.lua:62-69 loads): numbers are decimal tokens separated by blanks with a
    newline after each write; char/byte data (strings, Byte/CharStorage) is raw bytes followed by a newline."""

    _WS = b" \t\r\n"

    def _token(self) -> bytes:
        f = self.f
        c = f.read(1)
        while c and c in self._WS:
            c = f.read(1)
        if not c:
            raise EOFError("truncated Torch7 ascii file")
        out = bytearray()
        while c and c not in self._WS:
            out += c
            c = f.read(1)
        return bytes(out)

    def _eat_newline(self):
        c = self.f.read(1)
        if c and c != b"\n":
            self.f.seek(-1, 1)

    def i32(self):
        return int(self._token())

    i64 = i32

    def f64(self):
        return float(self._token())

    def string(self):
        n = self.i32()
        b = self._read(n)
        self._eat_newline()
        return b.decode("latin-1")

    def array(self, n: int, dt: np.dtype) -> np.ndarray:
        if n == 0:
            return np.zeros(0, dtype=dt)
        if dt.itemsize == 1:                       # raw, like the binary format
            arr = np.frombuffer(self._read(n), dtype=dt).copy()
            self._eat_newline()
            return arr
        # one line of n blank-separated numbers
        line = bytearray()
        while True:
            chunk = self.f.readline()
            if not chunk:
                raise EOFError("truncated Torch7 ascii file")
            line += chunk
            vals = line.split()
            if len(vals) >= n:
                break
        if len(vals) != n:
            raise ValueError("Torch7 ascii storage: element count mismatch")
        return np.array([float(v) for v in vals], dtype=np.float64).astype(dt)


class _Writer:
    def __init__(self, f: BinaryIO):
        self.f = f
        self.next_ref = 1
        self.memo: Dict[Any, int] = {}
        self.keep = []

    def i32(self, v):
        self.f.write(struct.pack("<i", v))

    def i64(self, v):
        self.f.write(struct.pack("<q", v))

    def f64(self, v):
        self.f.write(struct.pack("<d", v))

    def string(self, s: str):
        b = s.encode("latin-1")
        self.i32(len(b))
        self.f.write(b)

    def array(self, arr: np.ndarray):
        self.f.write(arr.tobytes())

    def _ref(self, o, kind: str = "") -> bool:
        """Write the reference index; True if the object was already written.  ``kind`` separates a
        tensor from its own storage, which may be the same Python object."""
        key = (kind, id(o))
        if key in self.memo:
            self.i32(self.memo[key])
            return True
        self.memo[key] = self.next_ref
        self.keep.append(o)
        self.i32(self.next_ref)
        self.next_ref += 1
        return False

    def obj(self, o):
        if o is None:
            self.i32(TYPE_NIL)
        elif isinstance(o, bool):
            self.i32(TYPE_BOOLEAN)
            self.i32(1 if o else 0)
        elif isinstance(o, (int, float)):
            self.i32(TYPE_NUMBER)
            self.f64(float(o))
        elif isinstance(o, str):
            self.i32(TYPE_STRING)
            self.string(o)
        elif isinstance(o, (list, tuple)):
            self.obj_table({i + 1: v for i, v in enumerate(o)}, o)
        elif isinstance(o, dict):
            self.obj_table(o, o)
        elif isinstance(o, torch.Tensor):
            self.tensor(o)
        elif isinstance(o, T7Object):
            self.i32(TYPE_TORCH)
            if self._ref(o):
                return
            self.string("V 1")
            self.string(o.classname)
            self.obj(o.fields)
        elif isinstance(o, T7Storage):
            self.storage(o.classname, o.array, o)
        else:
            raise TypeError(f"cannot serialise {type(o)} to Torch7")

    def obj_table(self, d: dict, identity):
        self.i32(TYPE_TABLE)
        if self._ref(identity):
            return
        self.i32(len(d))
        for k, v in d.items():
            self.obj(k)
            self.obj(v)

    def storage(self, cls: str, arr: np.ndarray, identity):
        self.i32(TYPE_TORCH)
        if self._ref(identity, "storage"):
            return
        self.string("V 1")
        self.string(cls)
        self.i64(arr.size)
        self.array(np.ascontiguousarray(arr).astype(_STORAGE_DTYPES[cls], copy=False))

    def tensor(self, t: torch.Tensor):
        orig = t
        t = t.detach().cpu()
        kind = _TORCH_TO_T7[t.dtype]
        self.i32(TYPE_TORCH)
        if self._ref(orig):
            return
        self.string("V 1")
        self.string(f"torch.{kind}Tensor")
        self.i32(t.dim())
        for s in t.shape:
            self.i64(s)
        tc = t.contiguous()
        for s in tc.stride():
            self.i64(s)
        self.i64(1)
        if tc.numel() == 0:
            self.i32(TYPE_NIL)
        else:
            self.storage(f"torch.{kind}Storage", tc.numpy().reshape(-1), tc)


class _AsciiWriter(_Writer):
    def i32(self, v):
        self.f.write(b"%d\n" % int(v))

    i64 = i32

    def f64(self, v):
        self.f.write(repr(float(v)).encode() + b"\n")

    def string(self, s: str):
        b = s.encode("latin-1")
        self.i32(len(b))
        self.f.write(b + b"\n")

    def array(self, arr: np.ndarray):
        if arr.size == 0:
            return
        if arr.dtype.itemsize == 1:
            self.f.write(arr.tobytes() + b"\n")
        elif arr.dtype.kind == "f":
            self.f.write(" ".join(repr(float(v)) for v in arr.tolist()).encode() + b"\n")
        else:
            self.f.write(" ".join(str(int(v)) for v in arr.tolist()).encode() + b"\n")


def load(path_or_file, mode: str = "binary") -> Any:
    """``torch.load(file [, 'ascii'])``."""
    R = _AsciiReader if mode == "ascii" else _Reader
    if hasattr(path_or_file, "read"):
        return R(path_or_file).obj()
    with open(path_or_file, "rb") as f:
        return R(f).obj()


def save(path_or_file, obj, mode: str = "binary") -> None:
    """``torch.save(file, obj [, 'ascii'])``."""
    W = _AsciiWriter if mode == "ascii" else _Writer
    if hasattr(path_or_file, "write"):
        W(path_or_file).obj(obj)
        return
    with open(path_or_file, "wb") as f:
        W(f).obj(obj)


def loads(data: bytes) -> Any:
    return _Reader(io.BytesIO(data)).obj()


def dumps(obj) -> bytes:
    b = io.BytesIO()
    _Writer(b).obj(obj)
    return b.getvalue()


def flat_float_header(n: int) -> bytes:
    """The bytes that precede the raw data of a 1-d ``torch.FloatTensor`` of ``n`` elements in a Torch7 binary file
    (tensor header + storage header).  Writing these and then ``n`` little-endian floats gives a file ``torch.load``
    reads as the flat tensor -- what BiCNN saves and ``-loadmodel`` loads (bicnn.lua:259-261,590-594).  Used by the
    sharded checkpoint writer, where every rank adds its own slice of the data."""
    buf = io.BytesIO()
    w = _Writer(buf)
    tensor_id, storage_id = object(), object()
    w.i32(TYPE_TORCH)
    w._ref(tensor_id)
    w.string("V 1")
    w.string("torch.FloatTensor")
    w.i32(1)
    w.i64(n)
    w.i64(1)            # stride
    w.i64(1)            # storage offset, 1-based
    w.i32(TYPE_TORCH)
    w._ref(storage_id, "storage")
    w.string("V 1")
    w.string("torch.FloatStorage")
    w.i64(n)
    return buf.getvalue()


def table_to_list(t: dict) -> list:
    """A 1..n keyed Lua table as a Python list."""
    return [t[i] for i in range(1, len(t) + 1)]


# ---- nn.Sequential of the CIFAR model (goot.lua:246-254 saves the whole model) --------------------------
def cifar_model_to_t7(theta: torch.Tensor) -> T7Object:
    """Build the Torch7 object graph of asyncsgd/models/7-layers-cnn.lua with parameters ``theta``
    (flat, getParameters() order) so that ``torch.load`` in Torch7 yields a usable ``nn.Sequential``."""
    theta = theta.detach().float().cpu()
    o = 0

    def take(*shape):
        nonlocal o
        n = int(np.prod(shape))
        t = theta[o:o + n].clone().view(*shape)
        o += n
        return t

    def conv(cin, cout, k):
        w, b = take(cout, cin, k, k), take(cout)
        return T7Object("nn.SpatialConvolution", {
            "nInputPlane": cin, "nOutputPlane": cout, "kW": k, "kH": k, "dW": 1, "dH": 1, "padW": 0, "padH": 0,
            "weight": w, "bias": b, "gradWeight": torch.zeros_like(w), "gradBias": torch.zeros_like(b),
            "output": torch.empty(0), "gradInput": torch.empty(0), "train": True})

    def simple(cls, **f):
        f.update({"output": torch.empty(0), "gradInput": torch.empty(0), "train": True})
        return T7Object(cls, f)

    def pool():
        return simple("nn.SpatialMaxPooling", kW=2, kH=2, dW=2, dH=2, padW=0, padH=0, ceil_mode=False,
                      indices=torch.empty(0))

    def linear(i, n):
        w, b = take(n, i), take(n)
        return T7Object("nn.Linear", {"weight": w, "bias": b, "gradWeight": torch.zeros_like(w),
                                      "gradBias": torch.zeros_like(b), "output": torch.empty(0),
                                      "gradInput": torch.empty(0), "train": True})

    relu = lambda: simple("nn.ReLU", threshold=0, val=0, inplace=False)
    mods = [conv(3, 64, 5), relu(), pool(), conv(64, 128, 5), relu(), pool(), conv(128, 64, 3), relu(),
            simple("nn.View", size=T7Storage("torch.LongStorage", np.array([256], dtype=np.int64)), numElements=256),
            linear(256, 256), relu(), simple("nn.Dropout", p=0.5, v2=True, noise=torch.empty(0)),
            linear(256, 10), simple("nn.LogSoftMax")]
    assert o == theta.numel(), "parameter vector does not match the 7-layer CNN"
    return T7Object("nn.Sequential", {"modules": mods, "output": torch.empty(0), "gradInput": torch.empty(0),
                                      "train": True})


def cifar_theta_from_t7(model: T7Object) -> torch.Tensor:
    """Inverse of :func:`cifar_model_to_t7`: flat parameters from a loaded ``nn.Sequential``."""
    mods = model.fields["modules"]
    mods = table_to_list(mods) if isinstance(mods, dict) else mods
    parts = []
    for m in mods:
        f = m.fields
        if isinstance(f, dict) and "weight" in f and f["weight"] is not None:
            parts += [f["weight"].reshape(-1).float(), f["bias"].reshape(-1).float()]
    return torch.cat(parts)
