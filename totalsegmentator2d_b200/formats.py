"""NIfTI-1 (``.nii`` / ``.nii.gz``) and MetaImage (``.mha`` / ``.mhd``) readers and writers.

The reference CLI accepts nrrd, nii, nii.gz, mha and mhd (reference ts2d/main.py:25) and reads all of them with
``sitk.ReadImage`` (reference ts2d/core/util/image.py:225-238).  SimpleITK is not part of this image, so the two
remaining formats are restated here the way ITK's NiftiImageIO / MetaImageIO present them to SimpleITK:

* ``array`` is ``[z][y][x]`` (x fastest), geometry is LPS (ITK's physical space).  NIfTI stores RAS: the first two rows of
  the affine are negated.  The affine comes from the qform when ``qform_code > 0``, else from the sform, else from
  ``pixdim`` alone; spacing = column norms, direction = normalised columns, origin = translation.
* ``scl_slope`` / ``scl_inter`` are applied when they are not the identity (ITK then reports a float image).
* MetaImage ``TransformMatrix`` lists the axis directions one after the other (= the transposed direction matrix).

Host-side I/O only: the hot path starts at the H2D copy of ``Image.array``.
"""
from __future__ import annotations

import gzip
import os
import struct
import zlib
from typing import Dict, Tuple

import numpy as np

from .image import Image

_NIFTI_DTYPES = {2: "u1", 4: "i2", 8: "i4", 16: "f4", 64: "f8", 256: "i1", 512: "u2", 768: "u4", 1024: "i8", 1280: "u8"}
_NIFTI_CODES = {v: k for k, v in _NIFTI_DTYPES.items()}


def _quat_to_mat(b: float, c: float, d: float, qfac: float) -> np.ndarray:
    a2 = 1.0 - (b * b + c * c + d * d)
    if a2 < 1e-7:                       # 180 degree rotation: renormalise (nifti1_io.c nifti_quatern_to_mat44)
        n = 1.0 / np.sqrt(b * b + c * c + d * d)
        b, c, d, a = b * n, c * n, d * n, 0.0
    else:
        a = np.sqrt(a2)
    R = np.array([[a * a + b * b - c * c - d * d, 2 * (b * c - a * d), 2 * (b * d + a * c)],
                  [2 * (b * c + a * d), a * a + c * c - b * b - d * d, 2 * (c * d - a * b)],
                  [2 * (b * d - a * c), 2 * (c * d + a * b), a * a + d * d - c * c - b * b]], dtype=np.float64)
    R[:, 2] *= qfac
    return R


def _sform_is_unsheared(M: np.ndarray, tol: float = 1e-3) -> bool:
    """True when the 3x3 part of an sform is rotation x per-axis scale: its normalised columns are mutually orthogonal"""
    n = np.linalg.norm(M, axis=0)
    if np.any(n < 1e-12) or not np.all(np.isfinite(M)):
        return False
    Q = M / n[None, :]
    return bool(np.abs(Q.T @ Q - np.eye(3)).max() < tol)


def read_nifti(path: str) -> Image:
    with open(path, "rb") as f:
        buf = f.read()
    if buf[:2] == b"\x1f\x8b":
        buf = gzip.decompress(buf)
    if len(buf) < 348:
        raise RuntimeError(f"{path}: not a NIfTI-1 file (shorter than its header)")
    end = "<"
    if struct.unpack("<i", buf[0:4])[0] != 348:
        end = ">"
        if struct.unpack(">i", buf[0:4])[0] != 348:
            raise RuntimeError(f"{path}: not a NIfTI-1 file (sizeof_hdr != 348)")
    magic = buf[344:348]
    if magic not in (b"n+1\0", b"ni1\0"):
        raise RuntimeError(f"{path}: unsupported NIfTI magic {magic!r}")
    dim = struct.unpack(end + "8h", buf[40:56])
    datatype, bitpix = struct.unpack(end + "2h", buf[70:74])
    pixdim = struct.unpack(end + "8f", buf[76:108])
    vox_offset, slope, inter = struct.unpack(end + "3f", buf[108:120])
    qform_code, sform_code = struct.unpack(end + "2h", buf[252:256])
    qb, qc, qd, qx, qy, qz = struct.unpack(end + "6f", buf[256:280])
    srow = np.array(struct.unpack(end + "12f", buf[280:328]), dtype=np.float64).reshape(3, 4)
    if datatype not in _NIFTI_DTYPES:
        raise RuntimeError(f"{path}: unsupported NIfTI datatype {datatype}")
    nd = int(dim[0])
    if not 1 <= nd <= 7:
        raise RuntimeError(f"{path}: bad dim[0] = {nd}")
    shape = [int(d) for d in dim[1:1 + nd]]
    while len(shape) > 3 and shape[-1] == 1:
        shape.pop()
    ncomp = 1
    if len(shape) == 5 and shape[3] == 1:          # NIfTI vector convention: dim[5] = components (intent vector)
        ncomp, shape = shape[4], shape[:3]
    if len(shape) > 3:
        raise RuntimeError(f"{path}: {len(shape)}-D NIfTI images are not supported")
    sd = len(shape)
    if magic == b"ni1\0":
        img_path = path[:-4] + ".img" if path.endswith(".hdr") else path
        with open(img_path, "rb") as f:
            data_buf, off = f.read(), 0
    else:
        data_buf, off = buf, int(vox_offset)
    dt = np.dtype(end + _NIFTI_DTYPES[datatype])
    count = int(np.prod(shape)) * ncomp
    data = np.frombuffer(data_buf, dtype=dt, count=count, offset=off)
    if ncomp > 1:
        arr = np.moveaxis(data.reshape([ncomp] + shape[::-1]), 0, -1)
    else:
        arr = data.reshape(shape[::-1])
    arr = np.ascontiguousarray(arr.astype(dt.newbyteorder("="), copy=False))
    if slope != 0.0 and not (slope == 1.0 and inter == 0.0) and np.isfinite(slope) and np.isfinite(inter):
        arr = (arr.astype(np.float64) * float(slope) + float(inter)).astype(np.float32)
    # affine voxel -> RAS mm.  Which transform wins when both are present follows ITK 5.3+ (SimpleITK 2.5.x, the reference's pin in
    # pyproject.toml): an sform with sform_code > 0 whose 3x3 part is a rotation times a per-axis scale (no shear) is preferred,
    # otherwise the qform, otherwise pixdim alone.  ITK 4 preferred the qform.  Spacing is pixdim (as ITK reads it) when positive.
    pd = np.array([abs(pixdim[1]) or 1.0, abs(pixdim[2]) or 1.0, abs(pixdim[3]) or 1.0])
    A = np.eye(4)
    if sform_code > 0 and _sform_is_unsheared(srow[:, :3]):
        A[:3, :] = srow
    elif qform_code > 0:
        qfac = -1.0 if pixdim[0] < 0 else 1.0
        R = _quat_to_mat(qb, qc, qd, qfac)
        A[:3, :3] = R * pd[None, :]
        A[:3, 3] = [qx, qy, qz]
    elif sform_code > 0:
        A[:3, :] = srow
    else:
        A[:3, :3] = np.diag(pd)
    A = np.diag([-1.0, -1.0, 1.0, 1.0]) @ A            # RAS -> LPS
    M = A[:3, :3]
    norms = np.linalg.norm(M, axis=0)
    norms[norms == 0] = 1.0
    D = M / norms[None, :]
    spacing = np.where(np.array(pixdim[1:4]) > 0, pd, norms)
    return Image(arr, tuple(spacing[:sd]), tuple(A[:sd, 3]), D[:sd, :sd].reshape(-1), is_vector=ncomp > 1)


def write_nifti(img: Image, path: str) -> None:
    """NIfTI-1 single file (n+1), qform and sform both set; gzip when the name ends in .gz."""
    arr = np.asarray(img.array)
    ncomp = 1
    if img.is_vector:
        # ITK's layout for vector pixels: a 5-D file, dim = (5, nx, ny, nz, 1, C), intent NIFTI_INTENT_VECTOR, component slowest
        ncomp = int(arr.shape[-1])
        arr = np.ascontiguousarray(np.moveaxis(arr, -1, 0))
    code = arr.dtype.str[1:]
    if code not in _NIFTI_CODES:
        raise RuntimeError(f"unsupported pixel type {arr.dtype} for NIfTI")
    sd = img.GetDimension()
    size = list(img.GetSize()) + [1] * (3 - sd)
    sp = list(img.GetSpacing()) + [1.0] * (3 - sd)
    D = np.eye(3)
    D[:sd, :sd] = img.direction_matrix()
    org = list(img.GetOrigin()) + [0.0] * (3 - sd)
    A = np.eye(4)
    A[:3, :3] = D * np.asarray(sp)[None, :]
    A[:3, 3] = org
    A = np.diag([-1.0, -1.0, 1.0, 1.0]) @ A            # LPS -> RAS
    R = A[:3, :3] / np.asarray(sp)[None, :]
    qfac = 1.0
    if np.linalg.det(R) < 0:
        R = R.copy()
        R[:, 2] *= -1
        qfac = -1.0
    tr = np.trace(R)                                    # rotation -> quaternion (nifti_mat44_to_quatern)
    a = 0.5 * np.sqrt(max(1.0 + tr, 0.0))
    if a > 0.5 * 1e-3 ** 0.5 and 1.0 + tr > 0.5:
        b, c, d = 0.25 * (R[2, 1] - R[1, 2]) / a, 0.25 * (R[0, 2] - R[2, 0]) / a, 0.25 * (R[1, 0] - R[0, 1]) / a
    else:
        xd, yd, zd = 1.0 + R[0, 0] - (R[1, 1] + R[2, 2]), 1.0 + R[1, 1] - (R[0, 0] + R[2, 2]), 1.0 + R[2, 2] - (R[0, 0] + R[1, 1])
        if xd > 1.0:
            b = 0.5 * np.sqrt(xd); c = 0.25 * (R[0, 1] + R[1, 0]) / b; d = 0.25 * (R[0, 2] + R[2, 0]) / b; a = 0.25 * (R[2, 1] - R[1, 2]) / b
        elif yd > 1.0:
            c = 0.5 * np.sqrt(yd); b = 0.25 * (R[0, 1] + R[1, 0]) / c; d = 0.25 * (R[1, 2] + R[2, 1]) / c; a = 0.25 * (R[0, 2] - R[2, 0]) / c
        else:
            d = 0.5 * np.sqrt(zd); b = 0.25 * (R[0, 2] + R[2, 0]) / d; c = 0.25 * (R[1, 2] + R[2, 1]) / d; a = 0.25 * (R[1, 0] - R[0, 1]) / d
        if a < 0:
            b, c, d = -b, -c, -d
    hdr = bytearray(352)
    struct.pack_into("<i", hdr, 0, 348)
    if ncomp > 1:
        struct.pack_into("<8h", hdr, 40, 5, size[0], size[1], size[2], 1, ncomp, 1, 1)
        struct.pack_into("<h", hdr, 68, 1007)           # intent_code = NIFTI_INTENT_VECTOR
    else:
        struct.pack_into("<8h", hdr, 40, 3, size[0], size[1], size[2], 1, 1, 1, 1)
    struct.pack_into("<2h", hdr, 70, _NIFTI_CODES[code], arr.dtype.itemsize * 8)
    struct.pack_into("<8f", hdr, 76, qfac, sp[0], sp[1], sp[2], 0, 1.0 if ncomp > 1 else 0, 0, 0)
    struct.pack_into("<3f", hdr, 108, 352.0, 1.0, 0.0)
    hdr[123] = 2                                        # xyzt_units: millimetres
    struct.pack_into("<2h", hdr, 252, 1, 1)
    struct.pack_into("<6f", hdr, 256, b, c, d, A[0, 3], A[1, 3], A[2, 3])
    struct.pack_into("<12f", hdr, 280, *A[:3, :].reshape(-1))
    hdr[344:348] = b"n+1\0"
    payload = bytes(hdr) + np.ascontiguousarray(arr.astype(arr.dtype.newbyteorder("<"), copy=False)).tobytes()
    if path.lower().endswith(".gz"):
        with gzip.open(path, "wb", compresslevel=1) as f:
            f.write(payload)
    else:
        with open(path, "wb") as f:
            f.write(payload)


_MET_TYPES = {"MET_CHAR": "i1", "MET_UCHAR": "u1", "MET_SHORT": "i2", "MET_USHORT": "u2", "MET_INT": "i4", "MET_UINT": "u4",
              "MET_LONG": "i4", "MET_ULONG": "u4", "MET_LONG_LONG": "i8", "MET_ULONG_LONG": "u8", "MET_FLOAT": "f4", "MET_DOUBLE": "f8"}
_MET_NAMES = {"i1": "MET_CHAR", "u1": "MET_UCHAR", "i2": "MET_SHORT", "u2": "MET_USHORT", "i4": "MET_INT", "u4": "MET_UINT",
              "i8": "MET_LONG_LONG", "u8": "MET_ULONG_LONG", "f4": "MET_FLOAT", "f8": "MET_DOUBLE"}


def _read_meta_header(buf: bytes) -> Tuple[Dict[str, str], int]:
    fields: Dict[str, str] = {}
    pos = 0
    while pos < len(buf):
        nl = buf.find(b"\n", pos)
        if nl == -1:
            nl = len(buf)
        line = buf[pos:nl].decode("latin-1").strip()
        pos = nl + 1
        if not line or "=" not in line:
            continue
        k, v = line.split("=", 1)
        fields[k.strip()] = v.strip()
        if k.strip() == "ElementDataFile":
            break
    return fields, pos


def read_metaimage(path: str) -> Image:
    with open(path, "rb") as f:
        buf = f.read()
    fields, off = _read_meta_header(buf)
    if "NDims" not in fields or "DimSize" not in fields or "ElementType" not in fields or "ElementDataFile" not in fields:
        raise RuntimeError(f"{path}: not a MetaImage file")
    nd = int(fields["NDims"])
    size = [int(v) for v in fields["DimSize"].split()]
    if nd not in (2, 3) or len(size) != nd:
        raise RuntimeError(f"{path}: {nd}-D MetaImage files are not supported")
    et = fields["ElementType"]
    if et not in _MET_TYPES:
        raise RuntimeError(f"{path}: unsupported ElementType {et}")
    ncomp = int(fields.get("ElementNumberOfChannels", "1"))
    msb = fields.get("BinaryDataByteOrderMSB", fields.get("ElementByteOrderMSB", "False")).lower() == "true"
    dt = np.dtype((">" if msb else "<") + _MET_TYPES[et])
    spacing = [float(v) for v in (fields.get("ElementSpacing") or fields.get("ElementSize") or " ".join(["1"] * nd)).split()]
    origin = [float(v) for v in (fields.get("Offset") or fields.get("Position") or fields.get("Origin") or " ".join(["0"] * nd)).split()]
    tm = fields.get("TransformMatrix") or fields.get("Orientation") or fields.get("Rotation")
    D = np.eye(nd) if tm is None else np.array([float(v) for v in tm.split()], dtype=np.float64).reshape(nd, nd).T
    src = fields["ElementDataFile"]
    if src == "LOCAL":
        data = buf[off:]
    elif src == "LIST" or "%" in src:
        raise RuntimeError(f"{path}: slice-list MetaImage files are not supported")
    else:
        with open(os.path.join(os.path.dirname(os.path.abspath(path)), src), "rb") as f:
            data = f.read()
        hs = int(fields.get("HeaderSize", "0"))
        if hs > 0:
            data = data[hs:]
    if fields.get("CompressedData", "False").lower() == "true":
        data = zlib.decompress(data)
    count = int(np.prod(size)) * ncomp
    if len(data) < count * dt.itemsize:
        raise RuntimeError(f"{path}: data section is shorter than DimSize says")
    arr = np.frombuffer(data, dtype=dt, count=count).reshape(size[::-1] + ([ncomp] if ncomp > 1 else []))
    arr = np.ascontiguousarray(arr.astype(dt.newbyteorder("="), copy=False))
    return Image(arr, spacing, origin, D.reshape(-1), is_vector=ncomp > 1)


def write_metaimage(img: Image, path: str, compress: bool = True) -> None:
    arr = np.ascontiguousarray(np.asarray(img.array))
    code = arr.dtype.str[1:]
    if code not in _MET_NAMES:
        raise RuntimeError(f"unsupported pixel type {arr.dtype} for MetaImage")
    nd = img.GetDimension()
    raw = arr.astype(arr.dtype.newbyteorder("<"), copy=False).tobytes()
    payload = zlib.compress(raw, 1) if compress else raw
    local = path.lower().endswith(".mha")
    data_name = "LOCAL" if local else os.path.basename(path)[:-4] + (".zraw" if compress else ".raw")
    fmt = lambda vs: " ".join(repr(float(v)) for v in vs)
    lines = ["ObjectType = Image", f"NDims = {nd}", "BinaryData = True", "BinaryDataByteOrderMSB = False",
             f"CompressedData = {'True' if compress else 'False'}"]
    if compress:
        lines.append(f"CompressedDataSize = {len(payload)}")
    lines += [f"TransformMatrix = {fmt(img.direction_matrix().T.reshape(-1))}", f"Offset = {fmt(img.GetOrigin())}",
              f"CenterOfRotation = {fmt([0.0] * nd)}", f"ElementSpacing = {fmt(img.GetSpacing())}",
              "DimSize = " + " ".join(str(int(s)) for s in img.GetSize())]
    if img.is_vector:
        lines.append(f"ElementNumberOfChannels = {img.GetNumberOfComponentsPerPixel()}")
    lines += [f"ElementType = {_MET_NAMES[code]}", f"ElementDataFile = {data_name}"]
    head = ("\n".join(lines) + "\n").encode("latin-1")
    if local:
        with open(path, "wb") as f:
            f.write(head + payload)
    else:
        with open(path, "wb") as f:
            f.write(head)
        with open(os.path.join(os.path.dirname(os.path.abspath(path)), data_name), "wb") as f:
            f.write(payload)


def read_any(path: str) -> Image:
    low = path.lower()
    if low.endswith(".nii") or low.endswith(".nii.gz"):
        return read_nifti(path)
    if low.endswith(".mha") or low.endswith(".mhd"):
        return read_metaimage(path)
    if low.endswith(".nrrd") or low.endswith(".nhdr"):
        from .nrrd import read_nrrd
        return read_nrrd(path)
    raise RuntimeError(f"unsupported image format: {path}")


def write_any(img: Image, path: str) -> None:
    low = path.lower()
    if low.endswith(".nii") or low.endswith(".nii.gz"):
        return write_nifti(img, path)
    if low.endswith(".mha") or low.endswith(".mhd"):
        return write_metaimage(img, path, True)
    if low.endswith(".nrrd"):
        from .nrrd import write_nrrd
        return write_nrrd(img, path, True)
    raise RuntimeError(f"unsupported image format: {path}")
