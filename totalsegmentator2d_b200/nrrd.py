"""Pure-Python NRRD reader/writer (the file format on both sides of the hot path).

The reference does all file I/O through ``sitk.ReadImage`` / ``sitk.WriteImage``
(reference ts2d/core/util/image.py:225-238, ts2d/tool.py:269-270).  SimpleITK is
not available here, so this module restates the subset of NRRD0004 that ITK's
NrrdImageIO emits and that the reference's assets use: scalar and vector
(``kinds: vector``) images, ``raw`` and ``gzip`` encodings, ``space directions``
with a leading ``none`` for the component axis, ``space`` or ``space dimension``,
and ``key:=value`` metadata lines (3D-Slicer ``SegmentN_*`` tags, reference
ts2d/core/util/meta.py:172-240).
"""
from __future__ import annotations

import gzip
import re
import zlib
from typing import Dict, List, Optional, Tuple

import numpy as np

from .image import Image

_TYPE_TO_DTYPE = {
    "signed char": "i1", "int8": "i1", "int8_t": "i1",
    "uchar": "u1", "unsigned char": "u1", "uint8": "u1", "uint8_t": "u1",
    "short": "i2", "short int": "i2", "signed short": "i2", "signed short int": "i2", "int16": "i2", "int16_t": "i2",
    "ushort": "u2", "unsigned short": "u2", "unsigned short int": "u2", "uint16": "u2", "uint16_t": "u2",
    "int": "i4", "signed int": "i4", "int32": "i4", "int32_t": "i4",
    "uint": "u4", "unsigned int": "u4", "uint32": "u4", "uint32_t": "u4",
    "longlong": "i8", "long long": "i8", "long long int": "i8", "signed long long": "i8", "int64": "i8", "int64_t": "i8",
    "ulonglong": "u8", "unsigned long long": "u8", "unsigned long long int": "u8", "uint64": "u8", "uint64_t": "u8",
    "float": "f4", "double": "f8",
}
_DTYPE_TO_TYPE = {
    "i1": "signed char", "u1": "unsigned char", "i2": "short", "u2": "unsigned short",
    "i4": "int", "u4": "unsigned int", "i8": "long long", "u8": "unsigned long long",
    "f4": "float", "f8": "double",
}
_VECTOR_KINDS = {"vector", "covariant-vector", "list", "point", "normal", "rgb-color", "rgba-color",
                 "3-color", "4-color", "complex", "2-vector", "3-vector", "4-vector", "3-gradient", "3-normal"}


def _parse_vec(tok: str) -> Optional[List[float]]:
    tok = tok.strip()
    if tok == "none":
        return None
    return [float(v) for v in tok.strip("()").split(",")]


def read_header(buf: bytes) -> Tuple[Dict[str, str], Dict[str, str], int]:
    """returns (fields, key/value metadata, data offset)"""
    if not buf.startswith(b"NRRD"):
        raise RuntimeError("not an NRRD file")
    end = buf.find(b"\n\n")
    crlf = buf.find(b"\r\n\r\n")
    if crlf != -1 and (end == -1 or crlf < end):
        head, off = buf[:crlf], crlf + 4
    elif end != -1:
        head, off = buf[:end], end + 2
    else:
        head, off = buf, len(buf)
    fields: Dict[str, str] = {}
    meta: Dict[str, str] = {}
    for line in head.decode("utf-8", errors="replace").splitlines()[1:]:
        line = line.rstrip("\r")
        if not line or line.startswith("#"):
            continue
        if ":=" in line:
            k, v = line.split(":=", 1)
            meta[k] = re.sub(r"\\(.)", lambda m: "\n" if m.group(1) == "n" else m.group(1), v)     # one left-to-right pass
        elif ":" in line:
            k, v = line.split(":", 1)
            fields[k.strip().lower()] = v.strip()
    return fields, meta, off


def read_nrrd(path: str) -> Image:
    with open(path, "rb") as f:
        buf = f.read()
    fields, meta, off = read_header(buf)
    if "data file" in fields or "datafile" in fields:
        raise RuntimeError("detached NRRD headers are not supported")
    dt = _TYPE_TO_DTYPE.get(fields["type"].lower())
    if dt is None:
        raise RuntimeError("unsupported NRRD type: {}".format(fields["type"]))
    endian = "<" if fields.get("endian", "little").lower() == "little" else ">"
    dtype = np.dtype(endian + dt) if dt[1] != "1" else np.dtype(dt)
    sizes = [int(s) for s in fields["sizes"].split()]
    ndim = int(fields["dimension"])
    assert ndim == len(sizes)
    enc = fields.get("encoding", "raw").lower()
    raw = buf[off:]
    if enc in ("gzip", "gz"):
        raw = zlib.decompress(raw, 16 + zlib.MAX_WBITS)
    elif enc != "raw":
        raise RuntimeError("unsupported NRRD encoding: {}".format(enc))
    n = int(np.prod(sizes))
    data = np.frombuffer(raw, dtype=dtype, count=n).astype(dtype.newbyteorder("="), copy=False)

    kinds = fields.get("kinds", "").split()
    dirs_raw = re.findall(r"none|\([^)]*\)", fields.get("space directions", ""))
    dirs = [_parse_vec(t) for t in dirs_raw] if dirs_raw else []
    # the component axis: ITK writes it first, with kind 'vector' and direction 'none'
    vec_axis = None
    for ax in range(ndim):
        k = kinds[ax] if ax < len(kinds) else "domain"
        d_none = (ax < len(dirs) and dirs[ax] is None)
        if k.lower() in _VECTOR_KINDS or (dirs and d_none):
            vec_axis = ax
            break
    arr = data.reshape(list(reversed(sizes)))   # NRRD axis 0 is fastest
    if vec_axis is not None:
        # numpy axis of NRRD axis a is ndim-1-a; move components last
        arr = np.moveaxis(arr, ndim - 1 - vec_axis, -1)
    sdim = ndim - (1 if vec_axis is not None else 0)

    space = fields.get("space", "").lower()
    sp_dirs = [d for d in dirs if d is not None] if dirs else []
    spacing = [1.0] * sdim
    direction = np.eye(sdim)
    if sp_dirs and len(sp_dirs) == sdim:
        m = np.asarray(sp_dirs, dtype=np.float64).T          # columns = image axes
        pdim = m.shape[0]
        if space in ("right-anterior-superior", "ras"):
            m[:2, :] *= -1.0
        elif space in ("left-anterior-superior", "las"):
            m[1, :] *= -1.0
        norms = np.linalg.norm(m, axis=0)
        norms[norms == 0] = 1.0
        spacing = [float(v) for v in norms]
        if pdim == sdim:
            direction = m / norms
    elif "spacings" in fields:
        sp = [s for s in fields["spacings"].split()]
        vals = [float(s) for i, s in enumerate(sp) if i != vec_axis and s.lower() != "nan"]
        if len(vals) == sdim:
            spacing = vals
    origin = [0.0] * sdim
    if "space origin" in fields:
        o = _parse_vec(fields["space origin"])
        if o is not None and len(o) == sdim:
            if space in ("right-anterior-superior", "ras"):
                o[0], o[1] = -o[0], -o[1]
            elif space in ("left-anterior-superior", "las"):
                o[1] = -o[1]
            origin = o
    img = Image(np.ascontiguousarray(arr), spacing, origin, direction.reshape(-1),
                is_vector=vec_axis is not None, meta=meta)
    return img


def _fmt(v: float) -> str:
    return repr(float(v)) if float(v) != int(v) or abs(v) > 1e15 else str(int(v))


def write_nrrd(img: Image, path: str, compress: bool = True, level: int = 1) -> None:
    """Writes what ITK's NrrdImageIO writes for the same image (LPS space, component axis first)."""
    arr = np.ascontiguousarray(img.array)
    dt = arr.dtype
    key = dt.str[1:]
    if key not in _DTYPE_TO_TYPE:
        if dt == np.bool_:
            arr, key = arr.astype(np.uint8), "u1"
        else:
            raise RuntimeError("unsupported pixel type for NRRD: {}".format(dt))
    sdim = img.GetDimension()
    size = list(img.GetSize())
    sizes = ([img.GetNumberOfComponentsPerPixel()] if img.is_vector else []) + size
    dmat = img.direction_matrix() * np.asarray(img.GetSpacing())[None, :]
    dir_tokens = ["(" + ",".join(_fmt(v) for v in dmat[:, a]) + ")" for a in range(sdim)]
    lines = ["NRRD0004",
             "# Complete NRRD file format specification at:",
             "# http://teem.sourceforge.net/nrrd/format.html",
             "type: " + _DTYPE_TO_TYPE[key],
             "dimension: {}".format(len(sizes))]
    if sdim == 3:
        lines.append("space: left-posterior-superior")
    else:
        lines.append("space dimension: {}".format(sdim))
    lines.append("sizes: " + " ".join(str(s) for s in sizes))
    lines.append("space directions: " + " ".join((["none"] if img.is_vector else []) + dir_tokens))
    lines.append("kinds: " + " ".join((["vector"] if img.is_vector else []) + ["domain"] * sdim))
    if arr.dtype.itemsize > 1:
        lines.append("endian: little")
    lines.append("encoding: " + ("gzip" if compress else "raw"))
    lines.append("space origin: (" + ",".join(_fmt(v) for v in img.GetOrigin()) + ")")
    for k in img.GetMetaDataKeys():
        v = img.GetMetaData(k).replace("\\", "\\\\").replace("\n", "\\n")
        lines.append("{}:={}".format(k, v))
    header = ("\n".join(lines) + "\n\n").encode("utf-8")
    payload = arr.astype(arr.dtype.newbyteorder("<"), copy=False).tobytes()
    if compress:
        payload = gzip.compress(payload, compresslevel=level, mtime=0)
    with open(path, "wb") as f:
        f.write(header)
        f.write(payload)
